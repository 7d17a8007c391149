// tracer_facade.cpp — drop-in replacement for the reference's src/tracer.cpp.
//
// Compile this file INSTEAD of src/tracer.cpp, against the reference's own headers
// (-I<reference>/src ...), and link libptb.so.  Everything else of the reference stays as it is:
// main.cpp, window.cpp, the scene loader (SceneFile.cpp, object.cpp), scene.cpp, shader.cpp's
// material table, camera.cpp.  The functions below have the signatures of src/tracer.h:16-31 and
// the meaning the callers rely on (src/main.cpp:94-145, src/window/window.cpp:74-77,254-294,454,
// 510,540); the work they used to do on 16 std::threads (src/tracer.cpp:113-185) is forwarded
// through the C ABI of include/ptb.h to the sm_100a wavefront kernels.
//
// What is read from the reference's globals, and when:
//   initTracer()  getTris()/getObjects()/getMaterials() -> ptb_upload_* (first call, or when the
//                 triangle count changed) + ptb_build_bvh / _subgrid / _subbeams; then the frame
//                 is zeroed and the per-pixel RNG reseeded x<<16|y (tracer.cpp:236-242)
//   trace()       per batch: getCamera() verbatim, material table (the UI edits both in place and
//                 calls reset()), getIntersectMode(), getNEE(), getPrimaryOnly(), debug flags
//   reset()       image + sample counts zeroed, RNG streams continue (tracer.cpp:195-217)
//
// Three knobs the reference keeps in file-static variables without getters are mirrored by small
// extension functions (declared in tracer_facade.h): grid settings, the max_depth cap of
// BASELINE.json's configs (absent in the reference), and the device index.
#include "tracer.h"

#include <atomic>
#include <chrono>
#include <climits>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <thread>
#include <vector>

#include "accelerationStructures/bvh.h"
#include "common.h"
#include "primitives/object.h"
#include "primitives/triangle.h"
#include "scene/scene.h"
#include "shader/shader.h"
#include "types/camera.h"
#include "types/texture.h"

#include "ptb.h"
#include "facade_internal.h"
#include "tracer_facade.h"

namespace {
int WIDTH = 800, HEIGHT = 600; // tracer.cpp:22-23
int MAX_SAMPLE_COUNT = INT_MAX;
bool finishedRendering = false, debugView = false, debugViewTris = false;
int debugScale = 10;
std::atomic<bool> running{true};
std::mutex mtx; // guards reinitTracer like tracer.cpp:162-173
bool reinitTracer = true;
bool clearPending = false, reseedPending = true;

ptb_group *grp = nullptr;  // one context per device (ptbFacadeSetDevices), tile-sharded; a single device by default
ptb_ctx *ctx = nullptr;    // member 0 (error text, structure sizes)
std::vector<int> deviceList{0};
size_t uploadedTris = (size_t)-1;
unsigned gridGs = 10, gridOs = 5, gridM = 40, beamGs = 10, beamM = 40; // main.cpp:42-44 defaults
bool gridBuilt = false, beamBuilt = false;
int maxDepth = 0;
bool useReferenceTree = false; // ptbFacadeUseReferenceTree
bool structuresFronted = false; // structures_facade.cpp is linked (bvh.h / lightFieldGrid*.h forward to the GPU builders too)
bool texturesUploaded = false;
int renderedSamples = 0;
std::vector<float> mirror; // host copy of the frame that tracerGetPixel serves (refreshed per batch)
uint64_t counters[PTB_CNT__N] = {0};

bool ok(int rc, const char *what) {
    if (rc == 0) return true;
    fprintf(stderr, "ptb facade: %s failed: %s\n", what, ptb_last_error(ctx)); // the reference prints and carries on
    return false;
}
bool gok(int rc, const char *what) { // calls on the whole group
    if (rc == 0) return true;
    fprintf(stderr, "ptb facade: %s failed: %s\n", what, ptb_group_last_error(grp));
    return false;
}
// the same per-context call on every member (scene and structures are replicated in each GPU's HBM)
template <typename F>
bool all(F f, const char *what) {
    ptb_ctx *first = ctx;
    bool good = true;
    for (int i = 0; good && i < ptb_group_size(grp); ++i) {
        ctx = ptb_group_ctx(grp, i); // ok() reports this member's error text
        good = ok(f(ctx), what);
    }
    ctx = first;
    return good;
}

bool ensureContext() {
    if (grp) return true;
    if (!gok(ptb_group_create(deviceList.data(), (int)deviceList.size(), &grp), "ptb_group_create")) return false;
    ctx = ptb_group_ctx(grp, 0);
    return true;
}

bool uploadMaterials() {
    auto &mats = *getMaterials();
    std::vector<ptbs_material> pm(mats.size());
    for (size_t i = 0; i < mats.size(); ++i) {
        const Material &m = mats[i];
        memset(&pm[i], 0, sizeof pm[i]);
        for (int k = 0; k < 3; ++k) pm[i].albedo[k] = m.pbr.albedo[k];
        pm[i].ior1 = m.pbr.refractiveIdx1;
        pm[i].ior2 = m.pbr.refractiveIdx2;
        pm[i].emission = m.pbr.emmision;
        pm[i].roughness = m.pbr.roughness;
        pm[i].w_lambert = m.weights.lambert;
        pm[i].w_reflection = m.weights.reflection;
        pm[i].w_refraction = m.weights.refraction;
        strncpy(pm[i].name, m.name.c_str(), sizeof pm[i].name - 1);
    }
    return all([&](ptb_ctx *c) { return ptb_upload_materials(c, pm.data(), (uint32_t)pm.size()); }, "ptb_upload_materials");
}

// Once per scene upload (the texel pools are rebuilt and gamma-expanded by the library); later material edits go
// through ptb_upload_materials alone, which keeps the bindings.
bool uploadTextures() {
    auto &mats = *getMaterials();
    // pbrProbs::texture / ::normal (shader.h:14-24): Texture::data is width*height Vector4 = 4 packed floats per texel
    std::vector<ptb_texture> tex;
    std::vector<int32_t> bind(2 * mats.size(), -1);
    for (size_t i = 0; i < mats.size(); ++i) {
        Texture *ts[2] = {&mats[i].pbr.texture, &mats[i].pbr.normal};
        for (int k = 0; k < 2; ++k) {
            Texture &t = *ts[k];
            if (t.data.empty() || t.width <= 0 || t.height <= 0 || t.data.size() != (size_t)t.width * t.height) continue;
            static_assert(sizeof(Vector4) == 4 * sizeof(float), "Vector4 must be 4 packed floats");
            bind[2 * i + k] = (int32_t)tex.size();
            tex.push_back(ptb_texture{&t.data[0].x, t.width, t.height});
        }
    }
    if (tex.empty() && !texturesUploaded) return true;
    texturesUploaded = !tex.empty();
    return all([&](ptb_ctx *c) { return ptb_upload_textures(c, tex.data(), (uint32_t)tex.size(), bind.data(), (uint32_t)mats.size()); }, "ptb_upload_textures");
}

bool uploadObjects() {
    auto &objs = getObjects();
    std::vector<ptbs_object> po(objs.size());
    for (size_t i = 0; i < objs.size(); ++i) {
        const Object &o = objs[i];
        memset(&po[i], 0, sizeof po[i]);
        po[i].start_idx = o.startIdx;
        po[i].end_idx = o.endIdx;
        po[i].surface_area = o.surfaceArea;
        po[i].n_materials = (uint32_t)std::min<size_t>(o.materials.size(), PTBS_MAX_OBJECT_MATERIALS);
        for (uint32_t k = 0; k < po[i].n_materials; ++k) po[i].materials[k] = o.materials[k];
        strncpy(po[i].name, o.name.c_str(), sizeof po[i].name - 1);
    }
    return all([&](ptb_ctx *c) { return ptb_upload_objects(c, po.data(), (uint32_t)po.size()); }, "ptb_upload_objects");
}

bool uploadReferenceTree();

// getTris() (280-byte AoS, triangle.h:7-14) -> the flat arrays of ptb_upload_triangles
bool uploadScene() {
    auto &tris = getTris();
    const size_t n = tris.size();
    std::vector<float> pos(n * 9), nrm(n * 9), uv(n * 6);
    std::vector<int32_t> mat(n);
    for (size_t i = 0; i < n; ++i) {
        Triangle &t = tris[i];
        for (int v = 0; v < 3; ++v) {
            for (int k = 0; k < 3; ++k) {
                pos[i * 9 + v * 3 + k] = t.vertices[v][k];
                nrm[i * 9 + v * 3 + k] = t.normal[v][k];
            }
            uv[i * 6 + v * 2] = t.uv[v].x;
            uv[i * 6 + v * 2 + 1] = t.uv[v].y;
        }
        mat[i] = t.materialIdx;
    }
    if (!all([&](ptb_ctx *c) { return ptb_upload_triangles(c, pos.data(), nrm.data(), uv.data(), mat.data(), n); }, "ptb_upload_triangles")) return false;
    if (!uploadMaterials() || !uploadTextures() || !uploadObjects()) return false;
    if (!all([&](ptb_ctx *c) { return ptb_build_bvh(c); }, "ptb_build_bvh")) return false; // constructBVH, scene.cpp:113-120
    gridBuilt = beamBuilt = false;
    uploadedTris = n;
    if (useReferenceTree && n > 0 && structuresFronted) {
        fprintf(stderr, "ptb facade: the reference tree was asked for, but bvh.h is fronted by structures_facade.cpp (no reference tree in this process)\n");
    } else if (useReferenceTree && n > 0) {
        if (!uploadReferenceTree() || !all([&](ptb_ctx *c) { return ptb_use_ref_bvh(c, 1); }, "ptb_use_ref_bvh")) return false;
    }
    return true;
}

// The reference's own BVH, read through the accessors it already exports (getNode / getNodeAABB /
// bvhGetTrisIndex, bvh.h:27-29; Object::root) after initScene built it (scene.cpp:113-120: the scene tree is
// the first constructBVH call, so its root is node 0).  With it on the device, mode BVH reproduces
// getIntersectionCount / getStructureIntersectionCount and the heat maps of the CPU tracer exactly.
bool uploadReferenceTree() {
    auto &objs = getObjects();
    std::vector<int32_t> roots(objs.size());
    for (size_t i = 0; i < objs.size(); ++i) roots[i] = objs[i].root;
    std::vector<ptb_ref_bvh_node> nodes;
    std::vector<uint32_t> idx;
    std::vector<int32_t> todo{0};
    std::vector<char> seen;
    while (!todo.empty()) {
        const int32_t i = todo.back();
        todo.pop_back();
        if (i < 0) continue;
        if ((size_t)i >= nodes.size()) {
            ptb_ref_bvh_node empty;
            memset(&empty, 0, sizeof empty);
            empty.child_left = empty.child_right = -1; // never reached: an empty leaf
            empty.has_tris = 1;
            nodes.resize((size_t)i + 1, empty);
            seen.resize((size_t)i + 1, 0);
        }
        if (seen[i]) continue;
        seen[i] = 1;
        const BvhNode &n = getNode((u32)i);
        const AABB &b = getNodeAABB((u32)n.AABBIdx);
        ptb_ref_bvh_node &o = nodes[i];
        for (int k = 0; k < 3; ++k) {
            o.bmin[k] = b.min[k];
            o.bmax[k] = b.max[k];
        }
        o.start_idx = n.startIdx;
        o.end_idx = n.endIdx;
        o.child_left = n.childLeft;
        o.child_right = n.childRight;
        o.has_tris = n.hasTris ? 1 : 0;
        if (n.childLeft == -1 && n.childRight == -1) {
            if (n.endIdx > (int)idx.size()) idx.resize((size_t)n.endIdx, 0u);
            for (int k = n.startIdx; k < n.endIdx; ++k) {
                idx[k] = bvhGetTrisIndex((u32)k);
                if (!n.hasTris && idx[k] < roots.size()) todo.push_back(roots[idx[k]]);
            }
        } else {
            todo.push_back(n.childLeft);
            todo.push_back(n.childRight);
        }
    }
    return all([&](ptb_ctx *c) {
        return ptb_upload_ref_bvh(c, nodes.data(), (uint32_t)nodes.size(), idx.data(), (uint32_t)idx.size(), roots.data(), (uint32_t)roots.size(), 0);
    }, "ptb_upload_ref_bvh");
}

bool ensureStructure(int mode) {
    if (mode == TWO_PLANE && !gridBuilt) { // constructGrid, scene.cpp:123
        // channels filled on each GPU (ptb_build_subgrid_gpu: same pools as the host builder, 58x faster on the Lucy stand-in)
        if (!all([&](ptb_ctx *c) { return ptb_build_subgrid_gpu(c, gridGs, gridOs, gridM); }, "ptb_build_subgrid_gpu")) return false;
        gridBuilt = true;
    }
    if (mode == SUB_BEAMS && !beamBuilt) { // buildGridsWithSubBeams (commented out at scene.cpp:125)
        // beam trees built level by level on each GPU (ptb_build_subbeams_gpu: same pools as the host's recursion)
        if (!all([&](ptb_ctx *c) { return ptb_build_subbeams_gpu(c, beamGs, beamM); }, "ptb_build_subbeams_gpu")) return false;
        beamBuilt = true;
    }
    return true;
}

bool pushSettings() {
    const Camera *cam = getCamera();
    const float o[3] = {cam->origin[0], cam->origin[1], cam->origin[2]}, f[3] = {cam->forward[0], cam->forward[1], cam->forward[2]};
    const float r[3] = {cam->right[0], cam->right[1], cam->right[2]}, u[3] = {cam->up[0], cam->up[1], cam->up[2]};
    if (!all([&](ptb_ctx *c) { return ptb_set_camera_basis(c, o, f, r, u, cam->focus, cam->dof, cam->fovScale); }, "ptb_set_camera_basis")) return false;
    if (!uploadMaterials() || !uploadObjects()) return false; // the UI edits materials in place (window.cpp:96-190); lights follow
    ptb_params p;
    memset(&p, 0, sizeof p);
    p.width = WIDTH;
    p.height = HEIGHT;
    p.mode = (int)getIntersectMode();
    p.nee = getNEE() ? 1 : 0;
    p.primary_only = getPrimaryOnly() ? 1 : 0;
    p.max_depth = maxDepth;
    p.reference_quirks = 1;
    p.debug_view = debugView ? (debugViewTris ? 2 : 1) : 0;
    p.debug_scale = debugScale;
    p.world = 1;
    p.tile_w = 32;
    p.tile_h = 8;
    p.count_tests = 1; // getIntersectionCount / getStructureIntersectionCount are part of tracer.h
    return gok(ptb_group_set_params(grp, &p), "ptb_group_set_params"); // rank / world per member
}

void refreshMirror() {
    mirror.resize((size_t)WIDTH * HEIGHT * 4);
    gok(ptb_group_read_image(grp, mirror.data()), "ptb_group_read_image"); // gathered from the owners on member 0
    ptb_group_read_counters(grp, counters);
}
} // namespace

// ---- shared with structures_facade.cpp (facade_internal.h) ----------------------------------------
namespace ptbfacade {
ptb_ctx *context() { return ensureContext() ? ctx : nullptr; }
bool sceneOnDevice() { return ensureContext() && (uploadedTris == getTris().size() || uploadScene()); }
bool structureOnDevice(int mode) { return sceneOnDevice() && ensureStructure(mode); }
void structuresAreFronted() { structuresFronted = true; }
} // namespace ptbfacade

// ---- tracer.h ------------------------------------------------------------------------------------
bool &getfinishedRendering() { return finishedRendering; }
bool &getDebugView() { return debugView; }
bool &getDebugShowTris() { return debugViewTris; }
int &getDebugScale() { return debugScale; }
int &getMaxSampleCount() { return MAX_SAMPLE_COUNT; }
unsigned long getIntersectionCount() { return (unsigned long)counters[PTB_CNT_TRI_TESTS]; }
unsigned long getStructureIntersectionCount() { return (unsigned long)counters[PTB_CNT_AS_TESTS]; }

Vector3 tracerGetPixel(int x, int y) {
    const size_t p = ((size_t)y * WIDTH + x) * 4;
    if (x < 0 || y < 0 || x >= WIDTH || y >= HEIGHT || p + 2 >= mirror.size()) return {0.0f, 0.0f, 0.0f};
    return {mirror[p], mirror[p + 1], mirror[p + 2]};
}

void setPixel(int x, int y, Vector3 &c) { // only the tracer itself called this in the reference (tracer.cpp:140,154)
    const size_t p = ((size_t)y * WIDTH + x) * 4;
    if (x < 0 || y < 0 || x >= WIDTH || y >= HEIGHT || p + 2 >= mirror.size()) return;
    mirror[p] = c[0];
    mirror[p + 1] = c[1];
    mirror[p + 2] = c[2];
}

void reset() { // tracer.cpp:195-217
    std::lock_guard<std::mutex> lock(mtx);
    std::fill(mirror.begin(), mirror.end(), 0.0f);
    resetScene();
    memset(counters, 0, sizeof counters);
    clearPending = true;
    renderedSamples = 0;
    reinitTracer = true;
}

void setWindowSize(int x, int y) {
    WIDTH = x;
    HEIGHT = y;
    mirror.assign((size_t)x * y * 4, 0.0f);
    reset();
}

Vector3 getWindowSize() { return {(float)WIDTH, (float)HEIGHT, 0.0f}; }

void initTracer() { // tracer.cpp:227-244
    std::lock_guard<std::mutex> lock(mtx);
    mirror.assign((size_t)WIDTH * HEIGHT * 4, 0.0f);
    memset(counters, 0, sizeof counters);
    reseedPending = true;
    renderedSamples = 0;
    reinitTracer = true;
}

void destroyTracer() { // tracer.cpp:245-251
    running = false;
    destroyScene();
}

// trace(testing), tracer.cpp:164-185: render until every pixel has MAX_SAMPLE_COUNT samples; with
// testing == false keep serving restarts (reset() from the UI) until destroyTracer().
void trace(bool testing) {
    bool run = true;
    while (run && running) {
        if (testing) run = false;
        bool clear = false, reseed = false;
        {
            std::lock_guard<std::mutex> lock(mtx);
            if (!reinitTracer) {
                if (!testing) std::this_thread::sleep_for(std::chrono::milliseconds(1));
                continue;
            }
            reinitTracer = false;
            clear = clearPending;
            reseed = reseedPending;
            clearPending = reseedPending = false;
        }
        finishedRendering = false;
        if (!ensureContext()) break;
        if (uploadedTris != getTris().size() && !uploadScene()) break;
        if (!ensureStructure((int)getIntersectMode()) || !pushSettings()) break;
        if (reseed) {
            if (!gok(ptb_group_reset(grp), "ptb_group_reset")) break;
        } else if (clear) {
            if (!gok(ptb_group_clear(grp), "ptb_group_clear")) break;
        }
        // progressive batches so the window sees the image converge and a reset() is honoured quickly
        int batch = testing ? MAX_SAMPLE_COUNT : 1;
        while (running && renderedSamples < MAX_SAMPLE_COUNT) {
            {
                std::lock_guard<std::mutex> lock(mtx);
                if (reinitTracer) break; // reset() arrived: restart with the new settings
            }
            const int todo = std::min(batch, MAX_SAMPLE_COUNT - renderedSamples);
            if (!gok(ptb_group_render(grp, todo), "ptb_group_render")) { // every GPU renders its tiles concurrently
                running = false;
                break;
            }
            renderedSamples += todo;
            refreshMirror();
            if (debugView) break; // the heat-map view takes one sample (tracer.cpp:136-142)
            if (batch < 64) batch *= 2;
        }
        finishedRendering = true;
    }
}

// ---- extensions (tracer_facade.h) ----------------------------------------------------------------
void ptbFacadeSetGridSettings(unsigned size, unsigned oSize, unsigned count) {
    gridGs = size;
    gridOs = oSize;
    gridM = count;
    gridBuilt = false;
}
void ptbFacadeSetGridBeamsSettings(unsigned size, unsigned count) {
    beamGs = size;
    beamM = count;
    beamBuilt = false;
}
void ptbFacadeSetMaxDepth(int depth) { maxDepth = depth; }
void ptbFacadeUseReferenceTree(bool on) {
    useReferenceTree = on;
    uploadedTris = (size_t)-1; // next trace() uploads again, with or without the tree
}
void ptbFacadeSetDevice(int device) { deviceList.assign(1, device); }
void ptbFacadeSetDevices(const int *devices, int n) { // before the first trace(): the frame is tile-sharded over them
    if (devices && n > 0) deviceList.assign(devices, devices + n);
}
unsigned long ptbFacadeStructureBytes(int mode) { return ctx ? (unsigned long)ptb_structure_bytes(ctx, mode) : 0ul; }
double ptbFacadeLastRenderMs() { return grp ? ptb_group_last_render_ms(grp) : 0.0; }
unsigned long ptbFacadeRaysTraced() { return (unsigned long)counters[PTB_CNT_RAYS]; }
