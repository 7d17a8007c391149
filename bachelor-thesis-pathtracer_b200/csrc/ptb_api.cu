// ptb_api.cu — host side of libptb.so: context, uploads, the wavefront scheduler and the
// extern "C" entry points declared in include/ptb.h.  No CPU rendering path exists here: every
// result comes from the kernels in ptb_kernels.cuh.
#include <cuda_runtime.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <exception>
#include <new>
#include <string>
#include <vector>

#include "ptb.h"
#include "ptb_bvh_build.h"
#include "ptb_grid_build.h"
#include "ptb_beam_gpu.h"
#include "ptb_grid_gpu.h"
#include "ptb_scene_loader.h"
#include "ptb_lbvh.h"
#include "ptb_kernels.cuh"
#include "ptb_mega.cuh"
// k_pool (warp pools) and k_duo (two paths per lane): measured and rejected in round 1.  Kept as
// scripts/ variants behind a build define (scripts/build_variants.py x=PTB_WITH_ALT_SCHEDULERS), not in the product library.
#ifdef PTB_WITH_ALT_SCHEDULERS
#include "ptb_pool.cuh"
#include "ptb_duo.cuh"
#else
#define PTB_POOL_MINBLOCKS 8
#define PTB_DUO_MINBLOCKS 8
#endif

#ifndef PTB_POOL_NP
#define PTB_POOL_NP 64 /* paths per warp pool */
#endif

using namespace ptb;

namespace {
const int kThreads = 128;      // CTA size of every traversal/logic kernel (__launch_bounds__)
const int kCtasPerSm = 8;      // persistent grid = SMs x this
const uint32_t kRing = 1u << 14; // per-iteration queue counters kept on the device (ring)
} // namespace

// a sub-wavefront: slots [base, base+n) with their own queues, counter rings, stream and events
struct Wave {
    cudaStream_t stream = nullptr;
    FrameDev F{};            // the frame with slot_base / n_slots of this wave
    uint32_t *dQueue[2] = {nullptr, nullptr}; // views into ctx->dQueue[*] at `base`
    uint32_t *dShadowQ = nullptr;
    uint32_t *dCnt = nullptr, *dShCnt = nullptr, *dCursor = nullptr; // [kRing] each
    uint32_t *hCnt = nullptr, *hShCnt = nullptr;                     // pinned mirrors
    std::vector<cudaEvent_t> evPool;                                 // 3 per iteration of a batch (profiling)
    cudaEvent_t evDone = nullptr;
    bool done = false;
    uint32_t carry = 0, carryShadow = 0;
};

struct WaveRange {
    Wave *b, *e;
    Wave *begin() const { return b; }
    Wave *end() const { return e; }
};

struct ptb_ctx {
    int device = 0;
    int sms = 148;
    int extendCtas = 12; // resident 128-thread CTAs per SM of the persistent BVH extend (40 registers/thread)
    int fetchKeep = 20;
    int useMega = 1;   // PTB_MEGA: 2 = warp-pool path kernel (ptb_pool.cuh), 1 = persistent-lane path kernel (ptb_mega.cuh), 0 = wavefront pipeline
    int poolCtas = PTB_POOL_MINBLOCKS;
    int duoCtas = PTB_DUO_MINBLOCKS; // PTB_DUO_CTAS
    int duoLogicKeep = 24, duoSampleKeep = 16, duoWalkMin = 20, duoRefillMin = 8, duoLeafKeep = 8; // PTB_DUO_LOGIC / _SAMPLE / _WALK / _REFILL
    void *dShScratch = nullptr; // NEE shadow records of the pool kernel (grid warps x NP)
    size_t shScratchN = 0;
    int megaCtas = PTB_MEGA_MINBLOCKS; // resident CTAs/SM of k_mega
    int logicKeep = -1; // PTB_LOGIC_KEEP: run a logic pass once this many lanes of a warp wait for it; -1 = by scene and mode (render_mega)
    int traceLeafFrac = 0; // PTB_TRACE_LEAF: the same switch for ptb_trace_rays (k_trace_bvh)
    int megaLeafFrac = -1; // PTB_MEGA_LEAF: k_mega leaves the node loop when fewer than walkers*this/32 lanes still descend; -1 = 16 for trees of >= 2^15 triangles, else off
    int leafMax = 3;    // PTB_LEAF_MAX: triangles per BVH leaf (<= 16)
    int megaChunk = 0;  // PTB_MEGA_CHUNK: samples per work item, 0 = chosen from pixels/lanes // PTB_LOGIC_KEEP: run a logic pass once this many lanes of a warp wait for it
    bool pendingRender = false, asyncRequested = false; // ptb_render_async: a path-kernel launch awaits ptb_sync
    bool matsChecked = false;   // every triangle's material index lies inside the material table (check_mode_ready)
    bool uniformSamples = true; // every owned pixel has the same sample count (false after ptb_write_state)
    int beamMaxDepth = 12; // guard on the reference's unbounded Subbeam recursion, ptb_set_build_limits  // refill a warp's idle lanes when fewer than this many are still walking
    cudaStream_t stream = nullptr;
    std::string err;
    bool poisoned = false;

    // host copy of the scene (what the loader handed over)
    std::vector<float> pos9, nrm9, uv6;
    std::vector<int32_t> mat;
    // textures (ptb_upload_textures): host copies + per-material bindings {albedo texture, normal map}
    std::vector<std::vector<float>> texData;
    std::vector<int32_t> texW, texH, matTex;
    bool haveTextures = false; // some material binds a texture: renders use the TEX build of the path kernel
    float2 *dUv = nullptr;
    TexDev *dTexTable = nullptr;
    float4 *dTexels = nullptr;
    std::vector<ptbs_object> objects;
    std::vector<ptbs_material> materials;
    std::vector<LightDev> lights;
    ptb_camera cam{};
    bool haveCam = false;
    bool camBasis = false; // camDev was given verbatim (ptb_set_camera_basis): skip the orthonormalisation
    CamDev camDev{};
    ptb_params prm{};
    bool haveParams = false;

    // structures
    BvhResult bvh;
    bool bvhBuilt = false, gridBuilt = false, beamBuilt = false;
    bool refBvhLoaded = false, useRefBvh = false; // ptb_upload_ref_bvh / ptb_use_ref_bvh
    float4 *dRefNode = nullptr;
    uint32_t *dRefIdx = nullptr;
    int *dRefObjRoot = nullptr;
    bool bvhOnGpu = false;   // built by ptb_build_bvh_gpu: no host copy of the nodes (c->bvh holds order + bounds only)
    uint64_t bvhNodes = 0;   // inner nodes of the tree in HBM
    double lastBuildMs = 0.0; // ptb_last_build_ms
    double lastBuildDeviceMs = 0.0; // device part of the last *_gpu build (CUDA events)
    std::vector<uint32_t> triOfId; // reference id -> leaf position
    uint64_t gridBytes = 0, beamBytes = 0;
    uint64_t refGridBytes = 0; // getMemory2Plane of the uploaded reference grids (ptb_upload_subgrid)
    uint64_t beamCounts[3] = {0, 0, 0}; // beams, leaf entries, root-table entries per axis (ptb_structure_digest)
    uint64_t gridCounts[3] = {0, 0, 0}; // grids, LUT cells, index entries (ptb_structure_digest)

    // device scene
    SceneDev S{};
    AttrDev A{};
    float4 *dTri = nullptr, *dNrm = nullptr, *dNodes = nullptr;
    float4 *dNodes4 = nullptr; // the 4-wide collapse of dNodes (k_collapse4), 128 B per BVH2 node; null unless PTB_WIDE
    int useWide = 0;           // PTB_WIDE: mode BVH renders walk the 4-wide collapse
    uint32_t *dTriOfId = nullptr, *dIdOfTri = nullptr;
    float *dVtx = nullptr;
    MatDev *dMats = nullptr;
    LightDev *dLights = nullptr;
    float4 *dGridHdr = nullptr;
    uint2 *dGridLut = nullptr;
    uint32_t *dGridIdx = nullptr;
    float4 *dBeam = nullptr;
    uint32_t *dBeamRoot[3] = {nullptr, nullptr, nullptr};
    uint32_t *dBeamIdx = nullptr;

    // frame + wavefront state
    FrameDev F{};
    int W = 0, H = 0;
    uint32_t nSlots = 0;
    uint32_t *dSlotPixel = nullptr;
    uint32_t *dQueue[2] = {nullptr, nullptr};
    uint32_t *dShadowQ = nullptr;
    uint32_t *dRings = nullptr; // maxWaves x 3 x kRing: per wave {main-queue sizes, shadow-queue sizes, extend work cursors} per iteration
    uint32_t *hRings = nullptr; // pinned mirror, maxWaves x 2 x kRing
    int maxWaves = 4;           // PTB_WAVES
    std::vector<Wave> allWaves; // maxWaves of them, created once
    int nWaves = 1;             // the first nWaves are in use (chosen per frame size in alloc_frame)
    unsigned long long *dCounters = nullptr;     // [8] paths, ref rays, tri tests, AS tests, no-extend queue entries
    int targetSamples = 0;
    uint64_t queueEntries = 0; // sum of main-queue sizes since reset
    uint64_t counters[PTB_CNT__N] = {0};
    cudaEvent_t evStart = nullptr, evStop = nullptr;
    double lastMs = 0.0;
    // per-stage device time (ptb_set_profiling): CUDA events around every k_extend / k_logic launch,
    // harvested at the scheduler's own batch syncs
    bool profile = false;
    double stageMs[2] = {0.0, 0.0};  // extend, logic (since reset)
    uint64_t stageLaunches[2] = {0, 0};
    bool frameDirty = true; // params changed: (re)allocate the frame on next reset

    // scratch for ptb_trace_rays / primary hits
    float *dScratchF = nullptr;
    int *dScratchI = nullptr;
    float *dRays = nullptr;
    uint64_t scratchN = 0;
};

static std::string g_createErr = "";
namespace {
int finish_mega(ptb_ctx *c);
}
// an enqueued ptb_render_async is completed before anything else touches the frame or the counters
#define SETTLE(c)                                   \
    do {                                            \
        if ((c)->pendingRender) {                   \
            const int rs_ = finish_mega(c);         \
            if (rs_) return rs_;                    \
        }                                           \
    } while (0)
static WaveRange active_waves(ptb_ctx *c) { return WaveRange{c->allWaves.data(), c->allWaves.data() + c->nWaves}; }

namespace {

int fail(ptb_ctx *c, int code, const std::string &msg) {
    if (c) c->err = msg;
    else g_createErr = msg;
    return code;
}

#define CK(call)                                                                                           \
    do {                                                                                                   \
        cudaError_t e_ = (call);                                                                           \
        if (e_ != cudaSuccess) {                                                                           \
            c->poisoned = true;                                                                            \
            return fail(c, PTB_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_));              \
        }                                                                                                  \
    } while (0)

#define GUARD(c)                                                                      \
    do {                                                                              \
        if (!(c)) return PTB_ERR_INVALID;                                             \
        if ((c)->poisoned) return fail((c), PTB_ERR_CUDA, "context poisoned by an earlier CUDA error: " + (c)->err); \
        cudaError_t e0_ = cudaSetDevice((c)->device);                                 \
        if (e0_ != cudaSuccess) return fail((c), PTB_ERR_CUDA, cudaGetErrorString(e0_)); \
    } while (0)

template <typename T>
void devFree(T *&p) {
    if (p) cudaFree(p);
    p = nullptr;
}

int grid_for(const ptb_ctx *c) { return c->sms * kCtasPerSm; }

void normalize3h(float *v) { // Eigen normalize (Dot.h:134-138), host float math, same order as the reference
    float z = v[0] * v[0] + (v[1] * v[1] + v[2] * v[2]);
    if (z > 0.0f) {
        float l = sqrtf(z);
        v[0] /= l;
        v[1] /= l;
        v[2] /= l;
    }
}
float dot3h(const float *a, const float *b) { return a[0] * b[0] + (a[1] * b[1] + a[2] * b[2]); }

// cameraSetForward -> orthoNormalized (camera.cpp:25-32, vector.cpp:22-27) with up=(0,1,0), right=(-1,0,0)
void make_camera(const ptb_camera &in, CamDev &out) {
    float f[3] = {in.forward[0], in.forward[1], in.forward[2]};
    normalize3h(f);
    float up[3] = {0, 1, 0}, right[3] = {-1, 0, 0};
    float d = dot3h(f, up);
    float b1[3] = {up[0] - d * f[0], up[1] - d * f[1], up[2] - d * f[2]};
    normalize3h(b1);
    float d0 = dot3h(f, right), d1 = dot3h(b1, right);
    float b2[3];
    for (int k = 0; k < 3; ++k) b2[k] = (right[k] - d0 * f[k]) - d1 * b1[k];
    normalize3h(b2);
    for (int k = 0; k < 3; ++k) {
        out.origin[k] = in.origin[k];
        out.forward[k] = f[k];
        out.up[k] = b1[k];
        out.right[k] = b2[k];
    }
    out.focus = in.focus;
    out.dof = in.dof;
    out.fovScale = tanf((in.fov_degrees * 3.14f / 180.0f) / 2.0f); // camera.cpp:16,43-45
}

ParamsDev make_params(const ptb_ctx *c) {
    ParamsDev P{};
    P.W = c->prm.width;
    P.H = c->prm.height;
    P.nee = c->prm.nee;
    P.primary_only = c->prm.primary_only;
    P.max_depth = c->prm.max_depth;
    P.quirks = c->prm.reference_quirks;
    P.debug_view = c->prm.debug_view;
    P.debug_scale = c->prm.debug_scale;
    P.target_samples = c->targetSamples;
    P.n_lights = (uint32_t)c->lights.size();
    if (c->camBasis) P.cam = c->camDev;
    else make_camera(c->cam, P.cam);
    return P;
}

} // namespace

// =============================================================================================
// lifetime
// =============================================================================================
extern "C" int ptb_create(int device, ptb_ctx **out) {
    if (!out) return PTB_ERR_INVALID;
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
        return fail(nullptr, PTB_ERR_CUDA, std::string("no CUDA device: libptb has no CPU fallback (") + cudaGetErrorString(e) + ")");
    if (device < 0 || device >= n) return fail(nullptr, PTB_ERR_INVALID, "device index out of range");
    ptb_ctx *c = new ptb_ctx();
    c->device = device;
    if ((e = cudaSetDevice(device)) != cudaSuccess) {
        delete c;
        return fail(nullptr, PTB_ERR_CUDA, cudaGetErrorString(e));
    }
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, device);
    c->sms = prop.multiProcessorCount;
    cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking);
    cudaEventCreate(&c->evStart);
    cudaEventCreate(&c->evStop);
    if (const char *e = getenv("PTB_MEGA")) c->useMega = std::max(0, std::min(3, atoi(e)));
#ifndef PTB_WITH_ALT_SCHEDULERS
    if (c->useMega > 1) {
        delete c;
        return fail(nullptr, PTB_ERR_INVALID, "PTB_MEGA=2/3 (k_pool / k_duo) are not in this build: compile with -DPTB_WITH_ALT_SCHEDULERS");
    }
#endif
    if (const char *e = getenv("PTB_DUO_CTAS")) c->duoCtas = std::max(1, std::min(PTB_DUO_MINBLOCKS, atoi(e)));
    if (const char *e = getenv("PTB_DUO_LOGIC")) c->duoLogicKeep = std::max(1, std::min(33, atoi(e)));
    if (const char *e = getenv("PTB_DUO_SAMPLE")) c->duoSampleKeep = std::max(1, std::min(33, atoi(e)));
    if (const char *e = getenv("PTB_DUO_WALK")) c->duoWalkMin = std::max(0, std::min(32, atoi(e)));
    if (const char *e = getenv("PTB_TRACE_LEAF")) c->traceLeafFrac = std::max(0, std::min(32, atoi(e)));
    if (const char *e = getenv("PTB_MEGA_LEAF")) c->megaLeafFrac = std::max(0, std::min(32, atoi(e)));
    if (const char *e = getenv("PTB_DUO_LEAF")) c->duoLeafKeep = std::max(1, std::min(33, atoi(e)));
    if (const char *e = getenv("PTB_DUO_REFILL")) c->duoRefillMin = std::max(1, std::min(64, atoi(e)));
    if (const char *e = getenv("PTB_POOL_CTAS")) c->poolCtas = std::max(1, std::min(16, atoi(e)));
    if (const char *e = getenv("PTB_MEGA_CTAS")) c->megaCtas = std::max(1, std::min(16, atoi(e)));
    if (const char *e = getenv("PTB_LEAF_MAX")) c->leafMax = std::max(1, std::min(16, atoi(e)));
    if (const char *e = getenv("PTB_MEGA_CHUNK")) c->megaChunk = std::max(0, atoi(e));
    if (const char *e = getenv("PTB_WIDE")) c->useWide = atoi(e) != 0;
    if (const char *e = getenv("PTB_LOGIC_KEEP")) c->logicKeep = std::max(1, std::min(64, atoi(e)));
    if (const char *e = getenv("PTB_WAVES")) c->maxWaves = std::max(1, std::min(16, atoi(e)));
    cudaMalloc(&c->dRings, (size_t)c->maxWaves * 3 * kRing * sizeof(uint32_t));
    cudaMallocHost(&c->hRings, (size_t)c->maxWaves * 2 * kRing * sizeof(uint32_t));
    c->allWaves.resize(c->maxWaves);
    for (int i = 0; i < c->maxWaves; ++i) {
        Wave &w = c->allWaves[i];
        cudaStreamCreateWithFlags(&w.stream, cudaStreamNonBlocking);
        cudaEventCreateWithFlags(&w.evDone, cudaEventDisableTiming);
        w.dCnt = c->dRings + (size_t)i * 3 * kRing;
        w.dShCnt = w.dCnt + kRing;
        w.dCursor = w.dShCnt + kRing;
        w.hCnt = c->hRings + (size_t)i * 2 * kRing;
        w.hShCnt = w.hCnt + kRing;
    }
    cudaMalloc(&c->dCounters, 16 * sizeof(unsigned long long));
    cudaMemset(c->dCounters, 0, 16 * sizeof(unsigned long long));
    if ((e = cudaGetLastError()) != cudaSuccess) {
        std::string m = cudaGetErrorString(e);
        delete c;
        return fail(nullptr, PTB_ERR_CUDA, m);
    }
    if (const char *e = getenv("PTB_EXTEND_CTAS")) c->extendCtas = std::max(1, std::min(16, atoi(e))); // tuning knobs
    if (const char *e = getenv("PTB_FETCH_KEEP")) c->fetchKeep = std::max(1, std::min(32, atoi(e)));
    c->cam.fov_degrees = 70.0f;
    c->cam.focus = 1.0f;
    c->cam.forward[2] = 1.0f;
    c->prm.reference_quirks = 1;
    c->prm.world = 1;
    c->prm.tile_w = 32;
    c->prm.tile_h = 8;
    *out = c;
    return PTB_OK;
}

static void free_frame(ptb_ctx *c) {
    devFree(c->F.image); devFree(c->F.samples); devFree(c->F.rng);
    devFree(c->dSlotPixel);
    devFree(c->F.ray_o); devFree(c->F.ray_d); devFree(c->F.hit); devFree(c->F.tp); devFree(c->F.light);
    devFree(c->F.sh_o); devFree(c->F.sh_d); devFree(c->F.sh_a); devFree(c->F.sh_b); devFree(c->F.sh_c); devFree(c->F.sh_hit);
    devFree(c->F.cnt_tri); devFree(c->F.cnt_as);
    devFree(c->dQueue[0]); devFree(c->dQueue[1]); devFree(c->dShadowQ);
}

extern "C" void ptb_destroy(ptb_ctx *c) {
    if (!c) return;
    cudaSetDevice(c->device);
    cudaDeviceSynchronize();
    free_frame(c);
    devFree(c->dNodes4);
    devFree(c->dTri); devFree(c->dNrm); devFree(c->dNodes); devFree(c->dTriOfId); devFree(c->dIdOfTri); devFree(c->dVtx); devFree(c->dNodes4);
    devFree(c->dMats); devFree(c->dLights);
    devFree(c->dGridHdr); devFree(c->dGridLut); devFree(c->dGridIdx);
    devFree(c->dBeam); devFree(c->dBeamIdx);
    devFree(c->dRefNode); devFree(c->dRefIdx); devFree(c->dRefObjRoot);
    devFree(c->dUv); devFree(c->dTexTable); devFree(c->dTexels);
    for (int k = 0; k < 3; ++k) devFree(c->dBeamRoot[k]);
    devFree(c->dRings); devFree(c->dCounters); devFree(c->dShScratch);
    devFree(c->dScratchF); devFree(c->dScratchI); devFree(c->dRays);
    if (c->hRings) cudaFreeHost(c->hRings);
    for (auto &w : c->allWaves) {
        for (auto &e : w.evPool) cudaEventDestroy(e);
        if (w.evDone) cudaEventDestroy(w.evDone);
        if (w.stream) cudaStreamDestroy(w.stream);
    }
    if (c->evStart) cudaEventDestroy(c->evStart);
    if (c->evStop) cudaEventDestroy(c->evStop);
    if (c->stream) cudaStreamDestroy(c->stream);
    delete c;
}

extern "C" const char *ptb_last_error(const ptb_ctx *c) { return c ? c->err.c_str() : g_createErr.c_str(); }

// =============================================================================================
// scene
// =============================================================================================
static int ptb_upload_triangles_impl(ptb_ctx *c, const float *pos9, const float *nrm9, const float *uv6, const int32_t *mat, uint64_t n) {
    GUARD(c);
    if (n && (!pos9 || !nrm9 || !mat)) return fail(c, PTB_ERR_INVALID, "null triangle arrays");
    if (n >= (1ull << 27)) return fail(c, PTB_ERR_INVALID, "more than 2^27 triangles");
    if (uv6) c->uv6.assign(uv6, uv6 + 6 * n); // Triangle::uv, read only by textured materials
    else c->uv6.clear();
    c->pos9.assign(pos9, pos9 + 9 * n);
    c->nrm9.assign(nrm9, nrm9 + 9 * n);
    c->mat.assign(mat, mat + n);
    c->matsChecked = false;
    c->bvhBuilt = c->gridBuilt = c->beamBuilt = false;
    c->refBvhLoaded = c->useRefBvh = false; // its indices refer to the previous triangle list
    return PTB_OK;
}

static int upload_lights(ptb_ctx *c) {
    // resetScene, scene.cpp:138-148: an object is a light if any of its materials has emission >= EPS
    c->lights.clear();
    for (auto &o : c->objects)
        for (uint32_t i = 0; i < o.n_materials && i < PTBS_MAX_OBJECT_MATERIALS; ++i) {
            uint32_t mi = o.materials[i];
            if (mi >= c->materials.size()) continue;
            if (c->materials[mi].emission < 0.00001f) continue;
            c->lights.push_back(LightDev{o.start_idx, o.end_idx, o.surface_area, 0});
            break;
        }
    devFree(c->dLights);
    if (!c->lights.empty()) {
        CK(cudaMalloc(&c->dLights, c->lights.size() * sizeof(LightDev)));
        CK(cudaMemcpy(c->dLights, c->lights.data(), c->lights.size() * sizeof(LightDev), cudaMemcpyHostToDevice));
    }
    c->A.lights = c->dLights;
    return PTB_OK;
}

static int ptb_upload_objects_impl(ptb_ctx *c, const ptbs_object *objects, uint32_t n) {
    GUARD(c);
    if (n && !objects) return fail(c, PTB_ERR_INVALID, "null objects");
    c->objects.assign(objects, objects + n);
    return upload_lights(c);
}

// MatDev table from the host material list + the texture bindings (if any)
static int upload_matdev(ptb_ctx *c) {
    const uint32_t n = (uint32_t)c->materials.size();
    std::vector<MatDev> md(n);
    const bool bound = c->matTex.size() == 2 * (size_t)n;
    c->haveTextures = false;
    for (uint32_t i = 0; i < n; ++i) {
        const ptbs_material &m = c->materials[i];
        MatDev &d = md[i];
        memset(&d, 0, sizeof d);
        for (int k = 0; k < 3; ++k) {
            d.albedo[k] = m.albedo[k];
            d.gamma[k] = powf(m.albedo[k], 2.2f); // gammaCorrect, shader.cpp:39-44 (host libm == the reference's)
        }
        d.ior1 = m.ior1; d.ior2 = m.ior2; d.emission = m.emission; d.roughness = m.roughness;
        d.wl = m.w_lambert; d.wr = m.w_reflection; d.wt = m.w_refraction;
        d.tex = d.nrm = -1;
        if (bound) { // texture t lives twice in the pool: 2t = gamma-expanded copy (albedo use), 2t+1 = raw copy (normal maps)
            const int32_t ta = c->matTex[2 * i], tn = c->matTex[2 * i + 1];
            if (ta >= 0) d.tex = 2 * ta;
            if (tn >= 0) d.nrm = 2 * tn + 1;
            if (ta >= 0 || tn >= 0) c->haveTextures = true;
        }
    }
    devFree(c->dMats);
    if (n) {
        CK(cudaMalloc(&c->dMats, n * sizeof(MatDev)));
        CK(cudaMemcpy(c->dMats, md.data(), n * sizeof(MatDev), cudaMemcpyHostToDevice));
    }
    c->A.mats = c->dMats;
    return PTB_OK;
}

static int ptb_upload_materials_impl(ptb_ctx *c, const ptbs_material *materials, uint32_t n) {
    GUARD(c);
    if (n && !materials) return fail(c, PTB_ERR_INVALID, "null materials");
    if (n != c->materials.size()) c->matTex.clear(); // another material list: the bindings no longer apply
    c->materials.assign(materials, materials + n);
    c->matsChecked = false;
    if (int r = upload_matdev(c)) return r;
    return upload_lights(c);
}

// Triangle::uv in leaf order, for the texture lookups (needs the BVH's order; no-op until both exist)
static int upload_uv(ptb_ctx *c) {
    devFree(c->dUv);
    c->A.uv = nullptr;
    const size_t n = c->mat.size();
    if (!c->haveTextures || !c->bvhBuilt || n == 0) return PTB_OK;
    std::vector<float2> uv(3 * n, make_float2(0.0f, 0.0f));
    if (c->uv6.size() == 6 * n)
        for (size_t k = 0; k < n; ++k) {
            const float *q = &c->uv6[6 * (size_t)c->bvh.order[k]];
            uv[3 * k] = make_float2(q[0], q[1]);
            uv[3 * k + 1] = make_float2(q[2], q[3]);
            uv[3 * k + 2] = make_float2(q[4], q[5]);
        }
    CK(cudaMalloc(&c->dUv, uv.size() * sizeof(float2)));
    CK(cudaMemcpy(c->dUv, uv.data(), uv.size() * sizeof(float2), cudaMemcpyHostToDevice));
    c->A.uv = c->dUv;
    return PTB_OK;
}

// Material textures: pbrProbs::texture / ::normal (shader.h:14-24), Texture = width x height float rgba
// (texture.h:7-11, byte/255.0f as loadTexture leaves them, texture.cpp:33-50).  The albedo copy is gamma-expanded
// here with the host libm — the reference applies powf(c, 2.2f) to every texel it reads (shader.cpp:214) — so the
// kernel never evaluates powf; normal maps are kept raw.
static int ptb_upload_textures_impl(ptb_ctx *c, const ptb_texture *textures, uint32_t n_textures, const int32_t *material_tex2, uint32_t n_materials) {
    GUARD(c);
    SETTLE(c);
    if (n_textures && !textures) return fail(c, PTB_ERR_INVALID, "null textures");
    if (n_materials != c->materials.size()) return fail(c, PTB_ERR_INVALID, "ptb_upload_materials first: one {albedo, normal} binding per material");
    if (n_materials && !material_tex2) return fail(c, PTB_ERR_INVALID, "null bindings");
    uint64_t texels = 0;
    for (uint32_t t = 0; t < n_textures; ++t) {
        if (!textures[t].rgba || textures[t].width <= 0 || textures[t].height <= 0 || textures[t].width > 32768 || textures[t].height > 32768)
            return fail(c, PTB_ERR_INVALID, "texture without data or with a bad size");
        texels += (uint64_t)textures[t].width * textures[t].height;
    }
    if (2 * texels >= (1ull << 32)) return fail(c, PTB_ERR_INVALID, "more than 2^31 texels");
    for (uint32_t i = 0; i < 2 * n_materials; ++i)
        if (material_tex2[i] < -1 || material_tex2[i] >= (int32_t)n_textures) return fail(c, PTB_ERR_INVALID, "texture binding out of range");
    c->texData.resize(n_textures);
    c->texW.resize(n_textures);
    c->texH.resize(n_textures);
    std::vector<TexDev> table(2 * (size_t)n_textures);
    std::vector<float4> pool(2 * texels);
    uint32_t at = 0;
    for (uint32_t t = 0; t < n_textures; ++t) {
        const size_t px = (size_t)textures[t].width * textures[t].height;
        c->texData[t].assign(textures[t].rgba, textures[t].rgba + 4 * px);
        c->texW[t] = textures[t].width;
        c->texH[t] = textures[t].height;
        const float *q = c->texData[t].data();
        table[2 * t] = TexDev{at, textures[t].width, textures[t].height, 0};
        for (size_t i = 0; i < px; ++i) pool[at + i] = make_float4(powf(q[4 * i], 2.2f), powf(q[4 * i + 1], 2.2f), powf(q[4 * i + 2], 2.2f), q[4 * i + 3]);
        at += (uint32_t)px;
        table[2 * t + 1] = TexDev{at, textures[t].width, textures[t].height, 0};
        for (size_t i = 0; i < px; ++i) pool[at + i] = make_float4(q[4 * i], q[4 * i + 1], q[4 * i + 2], q[4 * i + 3]);
        at += (uint32_t)px;
    }
    c->matTex.assign(material_tex2, material_tex2 + 2 * (size_t)n_materials);
    devFree(c->dTexTable); devFree(c->dTexels);
    if (n_textures == 0) { // unbind everything
        c->texData.clear();
        c->A.tex = nullptr;
        c->A.texels = nullptr;
        if (int r = upload_matdev(c)) return r;
        return upload_uv(c);
    }
    CK(cudaMalloc(&c->dTexTable, std::max<size_t>(1, table.size()) * sizeof(TexDev)));
    CK(cudaMalloc(&c->dTexels, std::max<size_t>(1, pool.size()) * sizeof(float4)));
    if (!table.empty()) CK(cudaMemcpy(c->dTexTable, table.data(), table.size() * sizeof(TexDev), cudaMemcpyHostToDevice));
    if (!pool.empty()) CK(cudaMemcpy(c->dTexels, pool.data(), pool.size() * sizeof(float4), cudaMemcpyHostToDevice));
    c->A.tex = c->dTexTable;
    c->A.texels = c->dTexels;
    if (int r = upload_matdev(c)) return r;
    return upload_uv(c);
}

extern "C" int ptb_set_camera(ptb_ctx *c, const ptb_camera *cam) {
    GUARD(c);
    if (!cam) return fail(c, PTB_ERR_INVALID, "null camera");
    c->cam = *cam;
    if (c->cam.fov_degrees == 0.0f) c->cam.fov_degrees = 70.0f;
    c->haveCam = true;
    c->camBasis = false;
    return PTB_OK;
}

extern "C" int ptb_set_camera_basis(ptb_ctx *c, const float origin[3], const float forward[3], const float right[3], const float up[3],
                                    float focus, float dof, float fov_scale) {
    GUARD(c);
    if (!origin || !forward || !right || !up) return fail(c, PTB_ERR_INVALID, "null camera vectors");
    for (int k = 0; k < 3; ++k) {
        c->camDev.origin[k] = origin[k];
        c->camDev.forward[k] = forward[k];
        c->camDev.right[k] = right[k];
        c->camDev.up[k] = up[k];
    }
    c->camDev.focus = focus;
    c->camDev.dof = dof;
    c->camDev.fovScale = fov_scale;
    c->haveCam = c->camBasis = true;
    return PTB_OK;
}

static int ptb_load_ptbs_impl(ptb_ctx *c, const char *path) {
    GUARD(c);
    FILE *f = path ? fopen(path, "rb") : nullptr;
    if (!f) return fail(c, PTB_ERR_IO, std::string("cannot open ") + (path ? path : "(null)"));
    ptbs_header h;
    if (fread(&h, sizeof h, 1, f) != 1 || h.magic != PTBS_MAGIC || (h.version != PTBS_VERSION && h.version != PTBS_VERSION_TEXTURES)) {
        fclose(f);
        return fail(c, PTB_ERR_IO, "not a .ptbs v1/v2 container");
    }
    // the header is untrusted: its counts must fit the file before anything is allocated from them
    size_t n = h.n_tris;
    {
        const long at = ftell(f);
        fseek(f, 0, SEEK_END);
        const uint64_t fileBytes = (uint64_t)ftell(f);
        fseek(f, at, SEEK_SET);
        const uint64_t need = (uint64_t)sizeof h + (uint64_t)h.n_tris * (9 + 9 + 6 + 1) * 4 + (uint64_t)h.n_objects * sizeof(ptbs_object) +
                              (uint64_t)h.n_materials * sizeof(ptbs_material);
        if (h.n_tris >= (1ull << 27) || h.n_objects > (1u << 24) || h.n_materials > (1u << 24) || need > fileBytes) {
            fclose(f);
            return fail(c, PTB_ERR_IO, "corrupt .ptbs header: element counts exceed the file size (or 2^27 triangles)");
        }
    }
    std::vector<float> pos(n * 9), nrm(n * 9), uv(n * 6);
    std::vector<int32_t> mat(n);
    std::vector<ptbs_object> objs(h.n_objects);
    std::vector<ptbs_material> mats(h.n_materials);
    bool ok = fread(pos.data(), 4, n * 9, f) == n * 9 && fread(nrm.data(), 4, n * 9, f) == n * 9 &&
              fread(uv.data(), 4, n * 6, f) == n * 6 && fread(mat.data(), 4, n, f) == n &&
              fread(objs.data(), sizeof(ptbs_object), objs.size(), f) == objs.size() &&
              fread(mats.data(), sizeof(ptbs_material), mats.size(), f) == mats.size();
    std::vector<int32_t> matTex(2 * mats.size(), -1);
    std::vector<std::vector<float>> texData;
    std::vector<ptb_texture> tex;
    if (ok && h.version >= PTBS_VERSION_TEXTURES && h.reserved[0] > 0) {
        ok = fread(matTex.data(), 4, matTex.size(), f) == matTex.size();
        for (uint32_t k = 0; ok && k < h.reserved[0]; ++k) {
            ptbs_texture_hdr th;
            ok = fread(&th, sizeof th, 1, f) == 1 && th.width > 0 && th.height > 0 && th.width <= 32768 && th.height <= 32768;
            if (!ok) break;
            texData.emplace_back((size_t)th.width * th.height * 4);
            ok = fread(texData.back().data(), 4, texData.back().size(), f) == texData.back().size();
            tex.push_back(ptb_texture{nullptr, th.width, th.height});
        }
    }
    fclose(f);
    if (!ok) return fail(c, PTB_ERR_IO, "truncated .ptbs container");
    int r;
    if ((r = ptb_upload_triangles(c, pos.data(), nrm.data(), uv.data(), mat.data(), n))) return r;
    if ((r = ptb_upload_materials(c, mats.data(), (uint32_t)mats.size()))) return r;
    if ((r = ptb_upload_objects(c, objs.data(), (uint32_t)objs.size()))) return r;
    for (size_t k = 0; k < tex.size(); ++k) tex[k].rgba = texData[k].data();
    // always re-bind: a scene without textures must not inherit the bindings of the scene this context held before
    if (tex.empty()) std::fill(matTex.begin(), matTex.end(), -1);
    if ((r = ptb_upload_textures(c, tex.data(), (uint32_t)tex.size(), matTex.data(), (uint32_t)mats.size()))) return r;
    ptb_camera cam{};
    for (int k = 0; k < 3; ++k) {
        cam.origin[k] = h.cam_pos[k];
        cam.forward[k] = h.cam_forward[k];
    }
    cam.focus = h.cam_focus;
    cam.dof = h.cam_dof;
    cam.fov_degrees = 70.0f;
    return ptb_set_camera(c, &cam);
}

// .scene / OBJ / MTL text -> the same uploads (csrc/ptb_scene_loader.cpp)
static int ptb_load_scene_file_impl(ptb_ctx *c, const char *scene_file, const char *base_dir, int flags) {
    GUARD(c);
    if (!scene_file) return fail(c, PTB_ERR_INVALID, "null scene file");
    LoadedScene s;
    std::string err;
    if (!load_scene_file(scene_file, base_dir ? base_dir : "", s, err, flags)) return fail(c, PTB_ERR_IO, err);
    int r;
    if ((r = ptb_upload_triangles(c, s.pos9.data(), s.nrm9.data(), s.uv6.data(), s.mat.data(), s.triangles()))) return r;
    if ((r = ptb_upload_materials(c, s.materials.data(), (uint32_t)s.materials.size()))) return r;
    if ((r = ptb_upload_objects(c, s.objects.data(), (uint32_t)s.objects.size()))) return r;
    {
        std::vector<ptb_texture> tex;
        for (const auto &t : s.texels) tex.push_back(ptb_texture{t.rgba.data(), t.width, t.height});
        std::vector<int32_t> bind(2 * s.materials.size(), -1); // no textures: clear what an earlier scene bound
        if (!tex.empty() && s.mat_tex.size() == bind.size()) bind = s.mat_tex;
        if ((r = ptb_upload_textures(c, tex.data(), (uint32_t)tex.size(), bind.data(), (uint32_t)s.materials.size()))) return r;
    }
    ptb_camera cam{};
    for (int k = 0; k < 3; ++k) {
        cam.origin[k] = s.cam_pos[k];
        cam.forward[k] = s.cam_forward[k];
    }
    cam.focus = s.cam_focus;
    cam.dof = s.cam_dof;
    cam.fov_degrees = 70.0f;
    return ptb_set_camera(c, &cam);
}

extern "C" int ptb_convert_scene_file(const char *scene_file, const char *base_dir, const char *out_ptbs, int flags, char *err, uint64_t err_len) {
    LoadedScene s;
    std::string why;
    bool ok = scene_file && out_ptbs;
    if (!ok) why = "null path";
    ok = ok && load_scene_file(scene_file, base_dir ? base_dir : "", s, why, flags) && write_ptbs(out_ptbs, s, why);
    if (err && err_len) {
        const size_t n = std::min<size_t>(why.size(), (size_t)err_len - 1);
        memcpy(err, why.data(), n);
        err[n] = 0;
    }
    return ok ? PTB_OK : PTB_ERR_IO;
}

// =============================================================================================
// structures
// =============================================================================================
// the 4-wide collapse of the tree in dNodes (BvhWalk4): one thread per BVH2 node, for host- and GPU-built trees alike.  Not
// built (mode BVH keeps walking the BVH2) when PTB_WIDE is off or the tree is deeper than the wide walk's stack allows.
static int build_wide_nodes(ptb_ctx *c, int depth) {
    devFree(c->dNodes4);
    c->S.nodes4 = nullptr;
    if (!c->useWide || c->bvhNodes == 0 || depth > PTB_WIDE_MAX_DEPTH) return PTB_OK;
    CK(cudaMalloc(&c->dNodes4, (size_t)c->bvhNodes * 8 * sizeof(float4)));
    k_collapse4<<<(int)std::min<uint64_t>(65535, (c->bvhNodes + 255) / 256), 256, 0, c->stream>>>(c->dNodes, c->dNodes4, (uint32_t)c->bvhNodes);
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(c->stream));
    c->S.nodes4 = c->dNodes4;
    return PTB_OK;
}

static int ptb_build_bvh_impl(ptb_ctx *c) {
    GUARD(c);
    const uint64_t n = c->mat.size();
    const auto t0 = std::chrono::steady_clock::now();
    build_bvh(c->pos9.data(), n, c->leafMax, c->bvh);
    c->lastBuildMs = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    // the walk pushes one deferred child per level: the tree must fit the per-lane stack
    if (c->bvh.max_depth >= PTB_STACK) return fail(c, PTB_ERR_INVALID, "BVH deeper than the traversal stack (PTB_STACK): degenerate input?");
    // leaf-ordered device arrays: positions as (v0, e1, e2) with the reference id in v0.w,
    // normals as (n0, n1, n2) with the material index in n0.w
    std::vector<float4> tri(3 * n), nrm(3 * n);
    c->triOfId.assign(n, 0);
    for (uint64_t k = 0; k < n; ++k) {
        const uint32_t id = c->bvh.order[k];
        c->triOfId[id] = (uint32_t)k;
        const float *p = &c->pos9[9 * (size_t)id], *q = &c->nrm9[9 * (size_t)id];
        float idBits, matBits;
        int32_t idi = (int32_t)id, mi = c->mat[id];
        memcpy(&idBits, &idi, 4);
        memcpy(&matBits, &mi, 4);
        tri[3 * k] = make_float4(p[0], p[1], p[2], idBits);
        tri[3 * k + 1] = make_float4(p[3] - p[0], p[4] - p[1], p[5] - p[2], 0.0f); // edge1 = v1 - v0 (triangle.cpp:10-11), fp32
        tri[3 * k + 2] = make_float4(p[6] - p[0], p[7] - p[1], p[8] - p[2], 0.0f); // edge2 = v2 - v0 (triangle.cpp:12-13)
        nrm[3 * k] = make_float4(q[0], q[1], q[2], matBits);
        nrm[3 * k + 1] = make_float4(q[3], q[4], q[5], 0.0f);
        nrm[3 * k + 2] = make_float4(q[6], q[7], q[8], 0.0f);
    }
    devFree(c->dTri); devFree(c->dNrm); devFree(c->dNodes); devFree(c->dTriOfId); devFree(c->dIdOfTri); devFree(c->dVtx);
    if (n) {
        CK(cudaMalloc(&c->dTri, 3 * n * sizeof(float4)));
        CK(cudaMalloc(&c->dNrm, 3 * n * sizeof(float4)));
        CK(cudaMalloc(&c->dTriOfId, n * sizeof(uint32_t)));
        CK(cudaMalloc(&c->dIdOfTri, n * sizeof(uint32_t)));
        CK(cudaMalloc(&c->dVtx, 9 * n * sizeof(float)));
        CK(cudaMemcpy(c->dTri, tri.data(), 3 * n * sizeof(float4), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(c->dNrm, nrm.data(), 3 * n * sizeof(float4), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(c->dTriOfId, c->triOfId.data(), n * sizeof(uint32_t), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(c->dIdOfTri, c->bvh.order.data(), n * sizeof(uint32_t), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(c->dVtx, c->pos9.data(), 9 * n * sizeof(float), cudaMemcpyHostToDevice));
    }
    const size_t nn = std::max<size_t>(1, c->bvh.nodes.size());
    CK(cudaMalloc(&c->dNodes, nn * sizeof(BvhNode64)));
    if (!c->bvh.nodes.empty())
        CK(cudaMemcpy(c->dNodes, c->bvh.nodes.data(), c->bvh.nodes.size() * sizeof(BvhNode64), cudaMemcpyHostToDevice));
    c->S.tri = c->dTri;
    c->S.nodes = c->dNodes;
    c->S.bvh_root = c->bvh.root;
    c->S.n_tris = (uint32_t)n;
    c->A.nrm = c->dNrm;
    c->A.tri_of_id = c->dTriOfId;
    c->S.ref_tri_pos = c->dTriOfId;
    c->A.id_of_tri = c->dIdOfTri;
    c->A.vtx = c->dVtx;
    c->bvhBuilt = true;
    c->bvhOnGpu = false;
    c->bvhNodes = c->bvh.nodes.size();
    c->gridBuilt = c->beamBuilt = false;
    if (int r = build_wide_nodes(c, c->bvh.max_depth)) return r;
    return upload_uv(c);
}

// constructBVH on the device (ptb_lbvh.cu): the triangles go up once in reference order, everything else —
// keys, sort, hierarchy, boxes, nodes, leaf-ordered arrays — is produced in HBM.  The only host copies
// kept are the leaf order (Subbeam builder input) and the scene bounds.
static int ptb_build_bvh_gpu_impl(ptb_ctx *c) {
    GUARD(c);
    SETTLE(c);
    const uint64_t n = c->mat.size();
    devFree(c->dTri); devFree(c->dNrm); devFree(c->dNodes); devFree(c->dTriOfId); devFree(c->dIdOfTri); devFree(c->dVtx);
    c->bvhBuilt = c->gridBuilt = c->beamBuilt = false;
    float *dNrm9 = nullptr;
    int32_t *dMat = nullptr;
    CK(cudaMalloc(&c->dVtx, std::max<size_t>(1, 9 * n) * sizeof(float)));
    cudaError_t e = cudaMalloc(&dNrm9, std::max<size_t>(1, 9 * n) * sizeof(float));
    if (e == cudaSuccess) e = cudaMalloc(&dMat, std::max<size_t>(1, n) * sizeof(int32_t));
    if (e == cudaSuccess && n) e = cudaMemcpyAsync(c->dVtx, c->pos9.data(), 9 * n * sizeof(float), cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess && n) e = cudaMemcpyAsync(dNrm9, c->nrm9.data(), 9 * n * sizeof(float), cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess && n) e = cudaMemcpyAsync(dMat, c->mat.data(), n * sizeof(int32_t), cudaMemcpyHostToDevice, c->stream);
    LbvhResult R;
    std::string msg;
    if (e == cudaSuccess) e = lbvh_build(c->dVtx, dNrm9, dMat, n, c->leafMax, PTB_STACK, c->stream, R, msg);
    else msg = std::string("staging the triangles: ") + cudaGetErrorString(e);
    cudaFree(dNrm9);
    cudaFree(dMat);
    if (e != cudaSuccess) {
        if (e != cudaErrorInvalidValue) c->poisoned = true;
        return fail(c, e == cudaErrorInvalidValue ? PTB_ERR_INVALID : PTB_ERR_CUDA, msg);
    }
    c->dTri = R.tri; c->dNrm = R.nrm; c->dNodes = R.nodes; c->dTriOfId = R.tri_of_id; c->dIdOfTri = R.id_of_tri;
    if (R.height >= PTB_STACK) {
        devFree(c->dTri); devFree(c->dNrm); devFree(c->dNodes); devFree(c->dTriOfId); devFree(c->dIdOfTri);
        return fail(c, PTB_ERR_INVALID, "BVH deeper than the traversal stack (PTB_STACK): degenerate input?");
    }
    c->bvh = BvhResult();
    c->bvh.root = R.root;
    c->bvh.max_depth = R.height;
    c->bvh.order.resize(n);
    if (n) CK(cudaMemcpy(c->bvh.order.data(), c->dIdOfTri, n * sizeof(uint32_t), cudaMemcpyDeviceToHost));
    for (int k = 0; k < 3; ++k) {
        c->bvh.lo[k] = R.lo[k];
        c->bvh.hi[k] = R.hi[k];
    }
    c->triOfId.clear();
    c->S.tri = c->dTri;
    c->S.nodes = c->dNodes;
    c->S.bvh_root = R.root;
    c->S.n_tris = (uint32_t)n;
    c->A.nrm = c->dNrm;
    c->A.tri_of_id = c->dTriOfId;
    c->S.ref_tri_pos = c->dTriOfId;
    c->A.id_of_tri = c->dIdOfTri;
    c->A.vtx = c->dVtx;
    c->bvhBuilt = c->bvhOnGpu = true;
    c->bvhNodes = R.n_nodes;
    c->lastBuildMs = c->lastBuildDeviceMs = R.ms;
    if (int r = build_wide_nodes(c, c->bvh.max_depth)) return r;
    return upload_uv(c);
}

extern "C" double ptb_last_build_ms(const ptb_ctx *c) { return c ? c->lastBuildMs : 0.0; }
extern "C" int ptb_bvh_depth(const ptb_ctx *c) { return (c && c->bvhBuilt) ? c->bvh.max_depth : -1; }

// The reference's own BVH — what constructBVH (bvh.cpp:222-390) left behind getNode / getNodeAABB / bvhGetTrisIndex
// (bvh.h:27-29) and Object::root — for count-exact traversal (refbvh_closest).  Validated here, because the kernel
// trusts it: every link in range, no cycles, and the worst-case stack of the reference's push order below PTB_STACK.
static int ptb_upload_ref_bvh_impl(ptb_ctx *c, const ptb_ref_bvh_node *nodes, uint32_t n_nodes, const uint32_t *indices, uint32_t n_indices,
                                  const int32_t *object_roots, uint32_t n_objects, int32_t scene_root) {
    GUARD(c);
    SETTLE(c);
    c->refBvhLoaded = c->useRefBvh = false;
    const uint64_t nTris = c->mat.size();
    if (!nodes || n_nodes == 0 || (n_indices && !indices) || (n_objects && !object_roots)) return fail(c, PTB_ERR_INVALID, "null reference-BVH arrays");
    if (scene_root < 0 || (uint32_t)scene_root >= n_nodes) return fail(c, PTB_ERR_INVALID, "scene_root out of range");
    for (uint32_t o = 0; o < n_objects; ++o)
        if (object_roots[o] < 0 || (uint32_t)object_roots[o] >= n_nodes) return fail(c, PTB_ERR_INVALID, "object root out of range");
    for (uint32_t i = 0; i < n_nodes; ++i) {
        const ptb_ref_bvh_node &nd = nodes[i];
        const bool leaf = nd.child_left == -1 && nd.child_right == -1;
        if (!leaf && (nd.child_left < 0 || nd.child_right < 0 || (uint32_t)nd.child_left >= n_nodes || (uint32_t)nd.child_right >= n_nodes))
            return fail(c, PTB_ERR_INVALID, "reference BVH: child index out of range");
        if (leaf) {
            if (nd.start_idx < 0 || nd.end_idx < nd.start_idx || (uint32_t)nd.end_idx > n_indices) return fail(c, PTB_ERR_INVALID, "reference BVH: leaf range outside the index array");
            for (int32_t k = nd.start_idx; k < nd.end_idx; ++k)
                if (indices[k] >= (nd.has_tris ? nTris : (uint64_t)n_objects)) return fail(c, PTB_ERR_INVALID, "reference BVH: leaf entry out of range");
        }
    }
    // worst-case stack entries while a subtree is being walked; -1 = being visited (cycle)
    std::vector<int> need(n_nodes, -2);
    struct Fr { int32_t node; int phase; };
    std::vector<Fr> st;
    auto children = [&](const ptb_ref_bvh_node &nd, std::vector<int32_t> &out) {
        out.clear();
        if (!(nd.child_left == -1 && nd.child_right == -1)) {
            out.push_back(nd.child_left);
            out.push_back(nd.child_right);
        } else if (!nd.has_tris)
            for (int32_t k = nd.start_idx; k < nd.end_idx; ++k) out.push_back(object_roots[indices[k]]);
    };
    std::vector<int32_t> ch;
    st.push_back({scene_root, 0});
    while (!st.empty()) {
        Fr &f = st.back();
        const int32_t node = f.node;
        if (f.phase == 0) {
            if (need[node] == -1) return fail(c, PTB_ERR_INVALID, "reference BVH: cycle");
            if (need[node] >= 0) {
                st.pop_back();
                continue;
            }
            need[node] = -1;
            f.phase = 1;
            children(nodes[node], ch);
            for (int32_t x : ch) {
                if (need[x] == -1) return fail(c, PTB_ERR_INVALID, "reference BVH: cycle");
                if (need[x] == -2) st.push_back({x, 0});
            }
            continue;
        }
        children(nodes[node], ch);
        int worst = 0;
        for (int32_t x : ch) worst = std::max(worst, need[x]);
        need[node] = ch.empty() ? 0 : (int)ch.size() - 1 + worst; // the siblings wait on the stack while one child is walked
        st.pop_back();
    }
    if (need[scene_root] + 1 >= PTB_STACK) return fail(c, PTB_ERR_INVALID, "reference BVH deeper than the traversal stack (PTB_STACK)");
    std::vector<float4> dev(3 * (size_t)n_nodes);
    for (uint32_t i = 0; i < n_nodes; ++i) {
        const ptb_ref_bvh_node &nd = nodes[i];
        float s, e, l, r, h;
        memcpy(&s, &nd.start_idx, 4); memcpy(&e, &nd.end_idx, 4); memcpy(&l, &nd.child_left, 4); memcpy(&r, &nd.child_right, 4);
        const int32_t ht = nd.has_tris ? 1 : 0;
        memcpy(&h, &ht, 4);
        dev[3 * (size_t)i] = make_float4(nd.bmin[0], nd.bmin[1], nd.bmin[2], s);
        dev[3 * (size_t)i + 1] = make_float4(nd.bmax[0], nd.bmax[1], nd.bmax[2], e);
        dev[3 * (size_t)i + 2] = make_float4(l, r, h, 0.0f);
    }
    devFree(c->dRefNode); devFree(c->dRefIdx); devFree(c->dRefObjRoot);
    CK(cudaMalloc(&c->dRefNode, dev.size() * sizeof(float4)));
    CK(cudaMalloc(&c->dRefIdx, std::max<size_t>(1, n_indices) * sizeof(uint32_t)));
    CK(cudaMalloc(&c->dRefObjRoot, std::max<size_t>(1, n_objects) * sizeof(int)));
    CK(cudaMemcpy(c->dRefNode, dev.data(), dev.size() * sizeof(float4), cudaMemcpyHostToDevice));
    if (n_indices) CK(cudaMemcpy(c->dRefIdx, indices, n_indices * sizeof(uint32_t), cudaMemcpyHostToDevice));
    if (n_objects) CK(cudaMemcpy(c->dRefObjRoot, object_roots, n_objects * sizeof(int), cudaMemcpyHostToDevice));
    c->S.ref_node = c->dRefNode;
    c->S.ref_idx = c->dRefIdx;
    c->S.ref_obj_root = c->dRefObjRoot;
    c->S.ref_root = scene_root;
    c->refBvhLoaded = true;
    return PTB_OK;
}

extern "C" int ptb_use_ref_bvh(ptb_ctx *c, int on) {
    GUARD(c);
    SETTLE(c);
    if (on && !c->refBvhLoaded) return fail(c, PTB_ERR_NOT_BUILT, "ptb_upload_ref_bvh first");
    c->useRefBvh = on != 0;
    return PTB_OK;
}

extern "C" int ptb_set_build_limits(ptb_ctx *c, int beam_max_depth) {
    GUARD(c);
    if (beam_max_depth < 0 || beam_max_depth > 32) return fail(c, PTB_ERR_INVALID, "beam_max_depth must be 0..32 (0 = unbounded like the reference)");
    c->beamMaxDepth = beam_max_depth;
    return PTB_OK;
}

extern "C" uint64_t ptb_structure_bytes(const ptb_ctx *c, int mode) {
    if (!c) return 0;
    const uint64_t tris = (uint64_t)c->mat.size() * 48ull;
    if (mode == PTB_MODE_BVH) return c->bvhBuilt ? c->bvhNodes * sizeof(BvhNode64) + tris : 0;
    if (mode == PTB_MODE_TWO_PLANE) return c->gridBuilt ? c->gridBytes + tris : 0;
    if (mode == PTB_MODE_SUB_BEAMS) return c->beamBuilt ? c->beamBytes + tris : 0;
    return tris;
}

// =============================================================================================
// params / frame
// =============================================================================================
extern "C" int ptb_set_params(ptb_ctx *c, const ptb_params *p) {
    GUARD(c);
    SETTLE(c);
    if (!p) return fail(c, PTB_ERR_INVALID, "null params");
    if (p->width <= 0 || p->height <= 0 || p->width > 65535 || p->height > 65535)
        return fail(c, PTB_ERR_INVALID, "width/height out of range (1..65535: the seed packs x<<16|y)");
    if (p->mode < 0 || p->mode > 3) return fail(c, PTB_ERR_INVALID, "mode must be 0..3");
    if (p->world < 1 || p->rank < 0 || p->rank >= p->world) return fail(c, PTB_ERR_INVALID, "bad rank/world");
    if (p->world > 1 && (p->tile_w <= 0 || p->tile_h <= 0)) return fail(c, PTB_ERR_INVALID, "bad tile size");
    const ptb_params &o = c->prm;
    if (!c->haveParams || o.width != p->width || o.height != p->height || o.rank != p->rank || o.world != p->world ||
        o.tile_w != p->tile_w || o.tile_h != p->tile_h || (p->count_tests || p->debug_view) != (o.count_tests || o.debug_view))
        c->frameDirty = true;
    c->prm = *p;
    c->haveParams = true;
    return PTB_OK;
}

static int alloc_frame(ptb_ctx *c) {
    free_frame(c);
    const ptb_params &p = c->prm;
    c->W = p.width;
    c->H = p.height;
    const size_t n = (size_t)c->W * c->H;
    // owned pixels: tile (tx,ty) belongs to rank (ty*tiles_x + tx) % world
    std::vector<uint32_t> sp;
    sp.reserve(n / (size_t)p.world + 1);
    if (p.world == 1) {
        sp.resize(n);
        for (size_t i = 0; i < n; ++i) sp[i] = (uint32_t)i;
    } else {
        const int tilesX = (c->W + p.tile_w - 1) / p.tile_w;
        for (int y = 0; y < c->H; ++y)
            for (int x = 0; x < c->W; ++x) {
                const int t = (y / p.tile_h) * tilesX + (x / p.tile_w);
                if (t % p.world == p.rank) sp.push_back((uint32_t)((size_t)y * c->W + x));
            }
    }
    c->nSlots = (uint32_t)sp.size();
    const size_t s = std::max<size_t>(1, c->nSlots);
    CK(cudaMalloc(&c->F.image, n * sizeof(float4)));
    CK(cudaMalloc(&c->F.samples, n * sizeof(int)));
    CK(cudaMalloc(&c->F.rng, n * sizeof(uint32_t)));
    CK(cudaMalloc(&c->dSlotPixel, s * sizeof(uint32_t)));
    CK(cudaMemcpy(c->dSlotPixel, sp.data(), c->nSlots * sizeof(uint32_t), cudaMemcpyHostToDevice));
    CK(cudaMalloc(&c->F.ray_o, s * sizeof(float4)));
    CK(cudaMalloc(&c->F.ray_d, s * sizeof(float4)));
    CK(cudaMalloc(&c->F.hit, s * sizeof(float4)));
    CK(cudaMalloc(&c->F.tp, s * sizeof(float4)));
    CK(cudaMalloc(&c->F.light, s * sizeof(float4)));
    CK(cudaMalloc(&c->F.sh_o, s * sizeof(float4)));
    CK(cudaMalloc(&c->F.sh_d, s * sizeof(float4)));
    CK(cudaMalloc(&c->F.sh_a, s * sizeof(float4)));
    CK(cudaMalloc(&c->F.sh_b, s * sizeof(float4)));
    CK(cudaMalloc(&c->F.sh_c, s * sizeof(float4)));
    CK(cudaMalloc(&c->F.sh_hit, s * sizeof(float4)));
    if (p.count_tests || p.debug_view) {
        CK(cudaMalloc(&c->F.cnt_tri, s * sizeof(uint32_t)));
        CK(cudaMalloc(&c->F.cnt_as, s * sizeof(uint32_t)));
    }
    CK(cudaMalloc(&c->dQueue[0], s * sizeof(uint32_t)));
    CK(cudaMalloc(&c->dQueue[1], s * sizeof(uint32_t)));
    CK(cudaMalloc(&c->dShadowQ, s * sizeof(uint32_t)));
    c->F.slot_pixel = c->dSlotPixel;
    c->F.n_slots = c->nSlots;
    c->F.slot_base = 0;
    // waves: contiguous slot ranges (multiples of 32 so warps never straddle), at least 128 Ki slots each
    int nw = (int)std::max<uint32_t>(1u, std::min<uint32_t>((uint32_t)c->maxWaves, c->nSlots / (128u << 10)));
    const uint32_t per = ((c->nSlots + nw - 1) / nw + 31u) & ~31u;
    c->nWaves = nw;
    for (int i = 0; i < nw; ++i) {
        Wave &w = c->allWaves[i];
        const uint32_t base = std::min<uint32_t>(c->nSlots, (uint32_t)i * per), end = std::min<uint32_t>(c->nSlots, base + per);
        w.F = c->F;
        w.F.slot_base = base;
        w.F.n_slots = end - base;
        w.dQueue[0] = c->dQueue[0] + base;
        w.dQueue[1] = c->dQueue[1] + base;
        w.dShadowQ = c->dShadowQ + base;
    }
    c->frameDirty = false;
    return PTB_OK;
}

static int reset_frame(ptb_ctx *c, int reseed);
extern "C" int ptb_reset(ptb_ctx *c) { return reset_frame(c, 1); }
extern "C" int ptb_clear(ptb_ctx *c) { return reset_frame(c, 0); }

static int reset_frame(ptb_ctx *c, int reseed) {
    GUARD(c);
    SETTLE(c);
    if (!c->haveParams) return fail(c, PTB_ERR_INVALID, "ptb_set_params first");
    if (c->frameDirty) {
        int r = alloc_frame(c);
        if (r) return r;
        reseed = 1; // a new frame has no streams to continue
    }
    k_reset<<<grid_for(c), 256, 0, c->stream>>>(c->F.image, c->F.samples, c->F.rng, c->W, c->H, reseed,
                                                     (uint32_t)c->prm.seed_offset * 0x9E3779B9u);
    CK(cudaMemsetAsync(c->dCounters, 0, 16 * sizeof(unsigned long long), c->stream));
    c->uniformSamples = true;
    CK(cudaStreamSynchronize(c->stream));
    c->targetSamples = 0;
    c->stageMs[0] = c->stageMs[1] = 0.0;
    c->stageLaunches[0] = c->stageLaunches[1] = 0;
    c->queueEntries = 0;
    memset(c->counters, 0, sizeof c->counters);
    c->counters[PTB_CNT_KERNELS] = 1;
    return PTB_OK;
}

static int check_mode_ready(ptb_ctx *c, int mode) {
    // every kernel reads A.mats[material id of the hit triangle]: an id outside the material table is rejected here, once per
    // (triangles, materials) pair, instead of becoming an out-of-bounds read on the device
    if (!c->matsChecked) {
        const int32_t nm = (int32_t)c->materials.size();
        for (size_t i = 0; i < c->mat.size(); ++i)
            if (c->mat[i] < 0 || c->mat[i] >= nm)
                return fail(c, PTB_ERR_INVALID, "triangle " + std::to_string(i) + " has material index " + std::to_string(c->mat[i]) +
                                                    " outside the " + std::to_string(nm) + " uploaded materials");
        c->matsChecked = true;
    }
    if (!c->bvhBuilt) return fail(c, PTB_ERR_NOT_BUILT, "ptb_build_bvh first (it also lays out the triangles)");
    if (mode == PTB_MODE_TWO_PLANE && !c->gridBuilt) return fail(c, PTB_ERR_NOT_BUILT, "ptb_build_subgrid first");
    if (mode == PTB_MODE_SUB_BEAMS && !c->beamBuilt) return fail(c, PTB_ERR_NOT_BUILT, "ptb_build_subbeams first");
    return PTB_OK;
}

// =============================================================================================
// the wavefront scheduler
// =============================================================================================
namespace {

// One wave = one contiguous range of slots with its own queues, counter rings and stream.  Waves are
// independent (a slot never leaves its wave), so their iterations overlap on the device: the end of
// wave A's extend launch — a few long rays on an otherwise empty GPU, ~0.03-0.1 ms per launch by
// measurement — is filled by wave B's CTAs instead of idling the machine.
template <int MODE, bool COUNT>
void launch_iteration(ptb_ctx *c, Wave &w, const ParamsDev &P, uint32_t it, bool nee, int evSlot) {
    // iteration `it`: extend rays of queue[it&1] (sizes dCnt[it], dShCnt[it]), then logic -> queue[(it+1)&1]
    const int g = grid_for(c);
    if (evSlot >= 0) cudaEventRecord(w.evPool[3 * evSlot], w.stream);
    if (MODE == 1)
        k_extend_bvh<COUNT><<<c->sms * c->extendCtas, kThreads, 0, w.stream>>>(c->S, w.F, w.dQueue[it & 1], w.dShadowQ, w.dCnt + it,
                                                                               nee ? w.dShCnt + it : nullptr, w.dCursor + it, c->fetchKeep);
    else
        k_extend<MODE, COUNT><<<g, kThreads, 0, w.stream>>>(c->S, w.F, w.dQueue[it & 1], w.dShadowQ, w.dCnt + it,
                                                            nee ? w.dShCnt + it : nullptr);
    if (evSlot >= 0) cudaEventRecord(w.evPool[3 * evSlot + 1], w.stream);
    k_logic<MODE, false, COUNT><<<g, kThreads, 0, w.stream>>>(P, c->S, c->A, w.F, w.dQueue[it & 1], w.dCnt + it, w.dQueue[(it + 1) & 1],
                                                              w.dCnt + it + 1, nee ? w.dShadowQ : nullptr,
                                                              nee ? w.dShCnt + it + 1 : nullptr, c->dCounters);
    if (evSlot >= 0) cudaEventRecord(w.evPool[3 * evSlot + 2], w.stream);
}

template <int MODE, bool COUNT>
int render_mode(ptb_ctx *c, int add) {
    c->targetSamples += add;
    const ParamsDev P = make_params(c);
    const bool nee = P.nee && P.n_lights > 0;
    const int g = grid_for(c);
    CK(cudaEventRecord(c->evStart, c->stream));
    for (auto &w : active_waves(c)) CK(cudaStreamWaitEvent(w.stream, c->evStart, 0));
    uint64_t kernels = 0, iterations = 0, rays = 0, shadowRays = 0;
    for (auto &w : active_waves(c)) {
        w.done = false;
        w.carry = w.carryShadow = 0;
    }
    bool first = true;
    int batch = 8;
    for (;;) {
        // one ring pass: counters [0..kRing) zeroed, slot 1 (or 0) seeded
        uint32_t it = first ? 1u : 0u;
        for (auto &w : active_waves(c)) {
            if (w.done) continue;
            CK(cudaMemsetAsync(w.dCnt, 0, kRing * sizeof(uint32_t), w.stream));
            CK(cudaMemsetAsync(w.dCursor, 0, kRing * sizeof(uint32_t), w.stream));
            if (nee) CK(cudaMemsetAsync(w.dShCnt, 0, kRing * sizeof(uint32_t), w.stream));
            if (first) {
                // START: every owned pixel that needs samples gets its primary ray -> queue[1], count dCnt[1]
                k_logic<MODE, true, COUNT><<<g, kThreads, 0, w.stream>>>(P, c->S, c->A, w.F, nullptr, nullptr, w.dQueue[1], w.dCnt + 1, nullptr,
                                                                         nullptr, c->dCounters);
                kernels++;
            } else {
                // continue a previous pass: the live queue is queue[0] by construction (see below)
                CK(cudaMemcpyAsync(w.dCnt, &w.carry, sizeof(uint32_t), cudaMemcpyHostToDevice, w.stream));
                if (nee) CK(cudaMemcpyAsync(w.dShCnt, &w.carryShadow, sizeof(uint32_t), cudaMemcpyHostToDevice, w.stream));
            }
        }
        first = false;
        bool allDone = false;
        uint32_t checked = it;
        while (!allDone && it + (uint32_t)batch + 2 < kRing) {
            const uint32_t it0 = it;
            for (int b = 0; b < batch; ++b, ++it)
                for (auto &w : active_waves(c))
                    if (!w.done) {
                        launch_iteration<MODE, COUNT>(c, w, P, it, nee, c->profile ? b : -1);
                        kernels += 2;
                    }
            for (auto &w : active_waves(c)) {
                if (w.done) continue;
                CK(cudaMemcpyAsync(w.hCnt + checked, w.dCnt + checked, (it + 1 - checked) * sizeof(uint32_t), cudaMemcpyDeviceToHost, w.stream));
                if (nee) CK(cudaMemcpyAsync(w.hShCnt + checked, w.dShCnt + checked, (it + 1 - checked) * sizeof(uint32_t), cudaMemcpyDeviceToHost, w.stream));
            }
            allDone = true;
            for (auto &w : active_waves(c)) {
                if (w.done) continue;
                CK(cudaStreamSynchronize(w.stream));
                if (c->profile)
                    for (uint32_t b = 0; b < it - it0; ++b) {
                        float e = 0.0f, l = 0.0f;
                        cudaEventElapsedTime(&e, w.evPool[3 * b], w.evPool[3 * b + 1]);
                        cudaEventElapsedTime(&l, w.evPool[3 * b + 1], w.evPool[3 * b + 2]);
                        c->stageMs[0] += e;
                        c->stageMs[1] += l;
                        c->stageLaunches[0]++;
                        c->stageLaunches[1]++;
                    }
                uint32_t live = 0;
                for (uint32_t k = checked; k < it; ++k) {
                    rays += w.hCnt[k];
                    if (nee) shadowRays += w.hShCnt[k];
                    live = std::max(live, w.hCnt[k]);
                }
                if (live && &w == c->allWaves.data()) iterations += it - checked;
                if (w.hCnt[it] == 0 && (!nee || w.hShCnt[it] == 0)) w.done = true;
                else allDone = false;
            }
            checked = it;
            if (batch < 64) batch *= 2;
        }
        if (allDone) break;
        // ring exhausted: make the live queue queue[0] for the next pass
        for (auto &w : active_waves(c)) {
            if (w.done) continue;
            w.carry = w.hCnt[it];
            if (it & 1) CK(cudaMemcpyAsync(w.dQueue[0], w.dQueue[1], (size_t)w.carry * sizeof(uint32_t), cudaMemcpyDeviceToDevice, w.stream));
            w.carryShadow = nee ? w.hShCnt[it] : 0u; // the single shadow queue buffer is position independent
            CK(cudaStreamSynchronize(w.stream));
        }
    }
    // everything the waves did is ordered before whatever follows on the context's stream
    for (auto &w : active_waves(c)) {
        CK(cudaEventRecord(w.evDone, w.stream));
        CK(cudaStreamWaitEvent(c->stream, w.evDone, 0));
    }
    CK(cudaEventRecord(c->evStop, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    float ms = 0.0f;
    cudaEventElapsedTime(&ms, c->evStart, c->evStop);
    c->lastMs = ms;
    unsigned long long dc[8];
    CK(cudaMemcpy(dc, c->dCounters, sizeof dc, cudaMemcpyDeviceToHost));
    c->queueEntries += rays;
    c->counters[PTB_CNT_SHADOW_RAYS] += shadowRays;
    c->counters[PTB_CNT_RAYS] = c->queueEntries - dc[4] + c->counters[PTB_CNT_SHADOW_RAYS];
    c->counters[PTB_CNT_ITERATIONS] += iterations;
    c->counters[PTB_CNT_KERNELS] += kernels;
    c->counters[PTB_CNT_PATHS] = dc[0];
    c->counters[PTB_CNT_REF_RAYS] = dc[1] + c->counters[PTB_CNT_SHADOW_RAYS];
    c->counters[PTB_CNT_TRI_TESTS] = dc[2];
    c->counters[PTB_CNT_AS_TESTS] = dc[3];
    return PTB_OK;
}

int finish_mega(ptb_ctx *c);

// One launch per render: the persistent-lane path kernel of ptb_mega.cuh.
template <int MODE, bool COUNT>
int render_mega(ptb_ctx *c, int add) {
    const int first = c->targetSamples;
    c->targetSamples += add;
    const ParamsDev P = make_params(c);
    // no more paths in flight than owned pixels: a pixel's samples are serial, so pixels bound the parallelism
    const bool pool = MODE < 4 && c->useMega == 2, duo = MODE < 4 && c->useMega == 3; // the reference tree and the wide tree run in k_mega only
    const uint64_t perCta = pool ? 4ull * PTB_POOL_NP : (duo ? 2ull * kThreads : (uint64_t)kThreads); // paths a CTA keeps in flight
    // Scarce pixels (fewer than two per lane of a full grid: small frames, 1/N of a frame per GPU): 8 resident CTAs per SM instead
    // of 12 and 16-sample items — profiles/r2_tile_share_sweep_bvh.jsonl, one GPU's share of an 8-GPU 1080p frame: 105.7 ms against
    // 114.1 ms (the rest of the gap to the ideal 75 ms is the spread of per-pixel cost: a pixel's samples are serial, and with ~1
    // pixel per lane the frame ends when its most expensive pixels do)
    int ctasPerSm = pool ? c->poolCtas : (duo ? c->duoCtas : std::min(c->megaCtas, PTB_MEGA_BLOCKS_FOR(MODE))); // never more than stay resident
    const bool scarce = !pool && !duo && (uint64_t)c->nSlots < 2ull * (uint64_t)c->sms * ctasPerSm * perCta && !getenv("PTB_MEGA_CTAS");
    if (scarce) ctasPerSm = std::min(ctasPerSm, 8);
    const int grid = (int)std::max<uint64_t>(1, std::min<uint64_t>((uint64_t)c->sms * ctasPerSm, ((uint64_t)c->nSlots + perCta - 1) / perCta));
    const uint64_t lanes = (uint64_t)grid * perCta;
    MegaArgs M{};
    // Work items are (pixel, chunk of samples); a lane whose pixel's previous chunk is still in flight parks
    // the item.  With many more pixels than lanes that never happens and 1-sample chunks end the frame within
    // one path of all lanes; when pixels are scarce (small frames, 1/N of a frame per GPU) longer chunks make
    // hand-offs rare at the price of a tail of `chunk` paths.  Non-uniform resumed state: one item per pixel.
    const uint64_t ratio = (uint64_t)c->nSlots / lanes;
    uint32_t chunk = ratio >= 2 ? 1u : (scarce ? 16u : 8u); // profiles/r1_sweep_mega_chunk.txt: 1-sample items keep neighbouring pixels in step (coherent primary rays)
    if (c->megaChunk > 0) chunk = (uint32_t)c->megaChunk;
    chunk = std::min<uint32_t>(chunk, (uint32_t)add);
    const uint32_t nChunks = ((uint32_t)add + chunk - 1) / chunk;
    const bool chunked = c->uniformSamples && (uint64_t)c->nSlots * (uint64_t)nChunks < (1ull << 32);
    // one item per pixel (non-uniform state): the item runs the pixel up to the target however far behind it is, like trace()
    // does until samples[x][y] reaches MAX_SAMPLE_COUNT (tracer.cpp:113-160)
    M.chunk_samples = chunked ? chunk : 0x7fffffffu;
    M.n_items = chunked ? c->nSlots * nChunks : c->nSlots;
    M.first_sample = first;
    M.check_ready = chunked ? 1 : 0;
    // profiles/r2_ab4_subgrid_knobs.txt: a small tree (Cornell) wants full shading passes (26 of 32 lanes); on deep structures a
    // traversal round is long, so waiting lanes cost more than a thinner shading pass: 20 (BVH), 18 (Subgrids: +5 %)
    const int autoKeep = c->mat.size() >= (1u << 15) ? (MODE == 2 ? 18 : 20) : 26;
    M.logic_keep = duo ? c->duoLogicKeep : (c->logicKeep > 0 ? c->logicKeep : autoKeep);
    M.sample_keep = c->duoSampleKeep;
    M.walk_min = c->duoWalkMin;
    M.refill_min = c->duoRefillMin;
    M.leaf_keep = duo ? c->duoLeafKeep : (c->megaLeafFrac >= 0 ? c->megaLeafFrac : (c->mat.size() >= (1u << 15) ? 16 : 0));
    M.cursor = c->dCounters + 8;
    M.counters = c->dCounters;
    CK(cudaMemsetAsync(c->dCounters + 8, 0, sizeof(unsigned long long), c->stream));
    CK(cudaEventRecord(c->evStart, c->stream));
    bool launched = false;
#ifdef PTB_WITH_ALT_SCHEDULERS
    if constexpr (MODE < 4) {
    if (pool) {
        launched = true;
        const bool nee = P.nee && P.n_lights > 0;
        const size_t need = nee ? (size_t)grid * 4 * PTB_POOL_NP : 0;
        if (need > c->shScratchN) {
            devFree(c->dShScratch);
            c->shScratchN = 0;
            CK(cudaMalloc(&c->dShScratch, need * sizeof(ShadowRec)));
            c->shScratchN = need;
        }
        const size_t smem = 4 * sizeof(PoolSmem<PTB_POOL_NP>);
        static bool attrSet[4][2] = {};
        if (!attrSet[MODE][COUNT ? 1 : 0]) {
            CK(cudaFuncSetAttribute(k_pool<MODE, COUNT, PTB_POOL_NP>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            CK(cudaFuncSetAttribute(k_pool<MODE, COUNT, PTB_POOL_NP>, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
            attrSet[MODE][COUNT ? 1 : 0] = true;
        }
        k_pool<MODE, COUNT, PTB_POOL_NP><<<grid, kThreads, smem, c->stream>>>(P, c->S, c->A, c->F, M, nee ? (ShadowRec *)c->dShScratch : nullptr);
    } else if (duo) {
        launched = true;
        if (P.nee && P.n_lights > 0) k_duo<MODE, COUNT, true><<<grid, kThreads, 0, c->stream>>>(P, c->S, c->A, c->F, M);
        else k_duo<MODE, COUNT, false><<<grid, kThreads, 0, c->stream>>>(P, c->S, c->A, c->F, M);
    }
    }
#endif
    if (!launched) {
        const bool nee = P.nee && P.n_lights > 0;
        const bool tex = c->haveTextures; // a material binds a texture: the build of the kernel with the texture branches
#define PTB_MEGA_LAUNCH(NEEV, EARLYV)                                                                                  \
    do {                                                                                                               \
        if (tex) k_mega<MODE, COUNT, NEEV, EARLYV, true><<<grid, kThreads, 0, c->stream>>>(P, c->S, c->A, c->F, M);     \
        else k_mega<MODE, COUNT, NEEV, EARLYV, false><<<grid, kThreads, 0, c->stream>>>(P, c->S, c->A, c->F, M);        \
    } while (0)
        if (PTB_MODE_HAS_ROUNDS(MODE) && M.leaf_keep > 0) { // deep structure: straggler exit in the descend loop
            if (nee) PTB_MEGA_LAUNCH(true, PTB_MODE_HAS_ROUNDS(MODE));
            else PTB_MEGA_LAUNCH(false, PTB_MODE_HAS_ROUNDS(MODE));
        } else if (nee) PTB_MEGA_LAUNCH(true, false);
        else PTB_MEGA_LAUNCH(false, false);
#undef PTB_MEGA_LAUNCH
    }
    CK(cudaEventRecord(c->evStop, c->stream));
    c->pendingRender = true;
    if (c->asyncRequested) return PTB_OK; // ptb_render_async: ptb_sync finishes the bookkeeping
    return finish_mega(c);
}

// waits for an enqueued k_mega / k_pool launch and folds its counters into the context
int finish_mega(ptb_ctx *c) {
    if (!c->pendingRender) return PTB_OK;
    c->pendingRender = false;
    CK(cudaStreamSynchronize(c->stream));
    float ms = 0.0f;
    cudaEventElapsedTime(&ms, c->evStart, c->evStop);
    c->lastMs = ms;
    c->stageMs[0] += ms; // one fused stage
    c->stageLaunches[0]++;
    unsigned long long dc[16];
    CK(cudaMemcpy(dc, c->dCounters, sizeof dc, cudaMemcpyDeviceToHost));
    c->counters[PTB_CNT_SHADOW_RAYS] = dc[6];
    c->counters[PTB_CNT_RAYS] = dc[5] + dc[6];
    c->counters[PTB_CNT_ITERATIONS] += 1;
    c->counters[PTB_CNT_KERNELS] += 1;
    c->counters[PTB_CNT_PATHS] = dc[0];
    c->counters[PTB_CNT_REF_RAYS] = dc[1] + dc[6];
    c->counters[PTB_CNT_TRI_TESTS] = dc[2];
    c->counters[PTB_CNT_AS_TESTS] = dc[3];
    return PTB_OK;
}

template <bool COUNT>
int render_dispatch(ptb_ctx *c, int add) {
    if (c->haveTextures && c->useMega != 1)
        return fail(c, PTB_ERR_INVALID, "textured materials are sampled by the default path kernel only (unset PTB_MEGA)");
    if (c->useRefBvh && c->prm.mode == PTB_MODE_BVH) {
        if (!c->useMega) return fail(c, PTB_ERR_INVALID, "the reference tree is traversed by the path kernel only (unset PTB_MEGA=0)");
        return render_mega<4, true>(c, add); // always the counting build: the counters are why this tree is here
    }
    if (c->useMega) switch (c->prm.mode) {
        case PTB_MODE_ALL: return render_mega<0, COUNT>(c, add);
        case PTB_MODE_BVH: return (c->S.nodes4 && c->useMega == 1) ? render_mega<5, COUNT>(c, add) : render_mega<1, COUNT>(c, add);
        case PTB_MODE_TWO_PLANE: return render_mega<2, COUNT>(c, add);
        default: return render_mega<3, COUNT>(c, add);
        }
    switch (c->prm.mode) {
    case PTB_MODE_ALL: return render_mode<0, COUNT>(c, add);
    case PTB_MODE_BVH: return render_mode<1, COUNT>(c, add);
    case PTB_MODE_TWO_PLANE: return render_mode<2, COUNT>(c, add);
    default: return render_mode<3, COUNT>(c, add);
    }
}

} // namespace

extern "C" int ptb_render(ptb_ctx *c, int add_samples) {
    GUARD(c);
    SETTLE(c);
    if (!c->haveParams) return fail(c, PTB_ERR_INVALID, "ptb_set_params first");
    if (add_samples <= 0) return fail(c, PTB_ERR_INVALID, "add_samples must be positive");
    int r = check_mode_ready(c, c->prm.mode);
    if (r) return r;
    if (c->frameDirty) {
        r = ptb_reset(c);
        if (r) return r;
    }
    if (c->nSlots == 0) return PTB_OK;
    const bool count = c->prm.count_tests || c->prm.debug_view;
    return count ? render_dispatch<true>(c, add_samples) : render_dispatch<false>(c, add_samples);
}
// Enqueues the render and returns: with the path kernels (one launch per render) nothing is read back until
// ptb_sync; the wavefront pipeline needs its queue sizes on the host and renders synchronously.
extern "C" int ptb_render_async(ptb_ctx *c, int add_samples) {
    GUARD(c);
    if (c->pendingRender) {
        int r = finish_mega(c);
        if (r) return r;
    }
    c->asyncRequested = true;
    const int r = ptb_render(c, add_samples);
    c->asyncRequested = false;
    return r;
}
extern "C" int ptb_sync(ptb_ctx *c) {
    GUARD(c);
    if (c->pendingRender) return finish_mega(c);
    CK(cudaStreamSynchronize(c->stream));
    return PTB_OK;
}
extern "C" double ptb_last_render_ms(const ptb_ctx *c) { return c ? c->lastMs : 0.0; }

extern "C" void *ptb_stream(ptb_ctx *c) { return c ? (void *)c->stream : nullptr; }

extern "C" int ptb_set_profiling(ptb_ctx *c, int on) {
    GUARD(c);
    if (on)
        for (auto &w : c->allWaves)
            if (w.evPool.empty()) {
                w.evPool.resize(3 * 64);
                for (auto &e : w.evPool) CK(cudaEventCreate(&e));
            }
    c->profile = on != 0;
    return PTB_OK;
}

extern "C" int ptb_read_stage_ms(ptb_ctx *c, double ms[2], uint64_t launches[2]) {
    GUARD(c);
    if (!ms || !launches) return fail(c, PTB_ERR_INVALID, "null");
    for (int k = 0; k < 2; ++k) {
        ms[k] = c->stageMs[k];
        launches[k] = c->stageLaunches[k];
    }
    return PTB_OK;
}

// =============================================================================================
// traversal-only entry points
// =============================================================================================
static int ensure_scratch(ptb_ctx *c, uint64_t n, bool rays) {
    if (n > c->scratchN) {
        devFree(c->dScratchF); devFree(c->dScratchI); devFree(c->dRays);
        c->scratchN = 0;
        CK(cudaMalloc(&c->dScratchF, n * sizeof(float)));
        CK(cudaMalloc(&c->dScratchI, 2 * n * sizeof(int)));
        c->scratchN = n;
    }
    if (rays && !c->dRays) CK(cudaMalloc(&c->dRays, c->scratchN * 6 * sizeof(float)));
    return PTB_OK;
}

extern "C" int ptb_primary_hits(ptb_ctx *c, float *t, int32_t *tri, int32_t *mat) {
    GUARD(c);
    SETTLE(c);
    if (!c->haveParams) return fail(c, PTB_ERR_INVALID, "ptb_set_params first");
    int r = check_mode_ready(c, c->prm.mode);
    if (r) return r;
    const uint64_t n = (uint64_t)c->prm.width * c->prm.height;
    if ((r = ensure_scratch(c, n, false))) return r;
    const ParamsDev P = make_params(c);
    const int g = grid_for(c);
    float *dT = c->dScratchF;
    int *dTri = c->dScratchI, *dMat = c->dScratchI + n;
    CK(cudaEventRecord(c->evStart, c->stream));
    switch (c->prm.mode) {
    case PTB_MODE_ALL: k_primary<0><<<g, kThreads, 0, c->stream>>>(P, c->S, c->A, dT, dTri, dMat); break;
    case PTB_MODE_BVH:
        if (c->useRefBvh) k_primary<4><<<g, kThreads, 0, c->stream>>>(P, c->S, c->A, dT, dTri, dMat);
        else k_primary<1><<<g, kThreads, 0, c->stream>>>(P, c->S, c->A, dT, dTri, dMat);
        break;
    case PTB_MODE_TWO_PLANE: k_primary<2><<<g, kThreads, 0, c->stream>>>(P, c->S, c->A, dT, dTri, dMat); break;
    default: k_primary<3><<<g, kThreads, 0, c->stream>>>(P, c->S, c->A, dT, dTri, dMat); break;
    }
    CK(cudaEventRecord(c->evStop, c->stream));
    c->counters[PTB_CNT_KERNELS]++;
    if (t) CK(cudaMemcpyAsync(t, dT, n * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
    if (tri) CK(cudaMemcpyAsync(tri, dTri, n * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    if (mat) CK(cudaMemcpyAsync(mat, dMat, n * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    float ms = 0;
    cudaEventElapsedTime(&ms, c->evStart, c->evStop);
    c->lastMs = ms;
    return PTB_OK;
}

extern "C" int ptb_trace_rays_device(ptb_ctx *c, const float *od6, uint64_t n, int mode, float *t, int32_t *tri) {
    GUARD(c);
    SETTLE(c);
    if (mode < 0 || mode > 3) return fail(c, PTB_ERR_INVALID, "mode must be 0..3");
    int r = check_mode_ready(c, mode);
    if (r) return r;
    if (n == 0) return PTB_OK;
    if (!od6 || !t || !tri) return fail(c, PTB_ERR_INVALID, "null ray / result pointers");
    const int g = (int)std::min<uint64_t>((uint64_t)grid_for(c), (n + 127) / 128);
    const bool count = c->prm.count_tests != 0;
    CK(cudaMemsetAsync(c->dCounters + 7, 0, sizeof(unsigned long long), c->stream));
    if (count) CK(cudaMemsetAsync(c->dCounters + 2, 0, 2 * sizeof(unsigned long long), c->stream)); // test counters = this call's
    CK(cudaEventRecord(c->evStart, c->stream));
#define PTB_TRACE(M)                                                                                                    \
    if (count) k_trace<M, true><<<g, 128, 0, c->stream>>>(c->S, od6, n, t, tri, c->dIdOfTri, c->dCounters);             \
    else k_trace<M, false><<<g, 128, 0, c->stream>>>(c->S, od6, n, t, tri, c->dIdOfTri, c->dCounters)
    switch (mode) {
    case PTB_MODE_ALL: PTB_TRACE(0); break;
    case PTB_MODE_BVH: {
        if (c->useRefBvh) { // count-exact walk of the reference's tree
            PTB_TRACE(4);
            break;
        }
        unsigned long long *cur = c->dCounters + 7; // slot 7 of the counter block: work cursor of this launch (zeroed above)
        const int gp = (int)std::min<uint64_t>((uint64_t)c->sms * c->extendCtas, (n + 127) / 128);
        const int leafFrac = c->traceLeafFrac; // off by default: coherent camera rays lose 10 % to it (config 3: 1533 -> 1383 Mrays/s)
        if (leafFrac > 0) {
            if (count) k_trace_bvh<true, true><<<gp, 128, 0, c->stream>>>(c->S, od6, n, t, tri, c->dIdOfTri, c->dCounters, cur, c->fetchKeep, leafFrac);
            else k_trace_bvh<false, true><<<gp, 128, 0, c->stream>>>(c->S, od6, n, t, tri, c->dIdOfTri, c->dCounters, cur, c->fetchKeep, leafFrac);
        } else if (count) k_trace_bvh<true, false><<<gp, 128, 0, c->stream>>>(c->S, od6, n, t, tri, c->dIdOfTri, c->dCounters, cur, c->fetchKeep, 0);
        else k_trace_bvh<false, false><<<gp, 128, 0, c->stream>>>(c->S, od6, n, t, tri, c->dIdOfTri, c->dCounters, cur, c->fetchKeep, 0);
        break;
    }
    case PTB_MODE_TWO_PLANE: PTB_TRACE(2); break;
    default: PTB_TRACE(3); break;
    }
#undef PTB_TRACE
    CK(cudaEventRecord(c->evStop, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    float ms = 0;
    cudaEventElapsedTime(&ms, c->evStart, c->evStop);
    c->lastMs = ms;
    c->counters[PTB_CNT_KERNELS]++;
    c->counters[PTB_CNT_RAYS] += n;
    if (count) {
        unsigned long long dc[8];
        CK(cudaMemcpy(dc, c->dCounters, sizeof dc, cudaMemcpyDeviceToHost));
        c->counters[PTB_CNT_TRI_TESTS] = dc[2];
        c->counters[PTB_CNT_AS_TESTS] = dc[3];
    }
    return PTB_OK;
}

extern "C" int ptb_trace_rays(ptb_ctx *c, const float *od6, uint64_t n, int mode, float *t, int32_t *tri) {
    GUARD(c);
    if (n == 0) return PTB_OK;
    if (!od6 || !t || !tri) return fail(c, PTB_ERR_INVALID, "null ray / result pointers");
    int r = ensure_scratch(c, n, true);
    if (r) return r;
    CK(cudaMemcpyAsync(c->dRays, od6, n * 6 * sizeof(float), cudaMemcpyHostToDevice, c->stream));
    if ((r = ptb_trace_rays_device(c, c->dRays, n, mode, c->dScratchF, c->dScratchI))) return r;
    CK(cudaMemcpyAsync(t, c->dScratchF, n * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
    CK(cudaMemcpyAsync(tri, c->dScratchI, n * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return PTB_OK;
}

// =============================================================================================
// read-back
// =============================================================================================
extern "C" int ptb_read_image(ptb_ctx *c, float *rgba) {
    GUARD(c);
    SETTLE(c);
    if (!rgba || !c->F.image) return fail(c, PTB_ERR_INVALID, "no frame / null buffer");
    CK(cudaMemcpyAsync(rgba, c->F.image, (size_t)c->W * c->H * sizeof(float4), cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return PTB_OK;
}

extern "C" int ptb_read_state(ptb_ctx *c, int32_t *samples, uint32_t *rng) {
    GUARD(c);
    SETTLE(c);
    if (!c->F.image) return fail(c, PTB_ERR_INVALID, "no frame");
    const size_t n = (size_t)c->W * c->H;
    if (samples) CK(cudaMemcpyAsync(samples, c->F.samples, n * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    if (rng) CK(cudaMemcpyAsync(rng, c->F.rng, n * sizeof(uint32_t), cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return PTB_OK;
}

extern "C" int ptb_write_state(ptb_ctx *c, const float *rgba, const int32_t *samples, const uint32_t *rng) {
    GUARD(c);
    SETTLE(c);
    if (!c->F.image) return fail(c, PTB_ERR_INVALID, "no frame (ptb_reset first)");
    if (!rgba || !samples || !rng) return fail(c, PTB_ERR_INVALID, "null state arrays");
    const size_t n = (size_t)c->W * c->H;
    CK(cudaMemcpyAsync(c->F.image, rgba, n * sizeof(float4), cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(c->F.samples, samples, n * sizeof(int), cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(c->F.rng, rng, n * sizeof(uint32_t), cudaMemcpyHostToDevice, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    int mx = 0;
    for (size_t i = 0; i < n; ++i) mx = std::max(mx, samples[i]);
    c->targetSamples = mx;
    c->uniformSamples = false;
    return PTB_OK;
}

// the reference's display expression (window.cpp:513-522) per channel, evaluated with the host libm
static unsigned char tone_byte_host(float v) {
    float p = powf(v, 0.4545f);
    p = fminf(p, 1.0f);
    return (unsigned char)(p * 255);
}
// t[b] = smallest float v in [0, 1] with tone_byte_host(v) >= b: binary search over the bit patterns (non-negative floats
// order like their bits); the step function is non-decreasing (checked exhaustively by tests/native/tonemap_check.cpp)
static const ToneTable &tone_thresholds() {
    static ToneTable T;
    static bool built = false;
    if (!built) {
        T.t[0] = 0.0f;
        for (int b = 1; b < 256; ++b) {
            uint32_t lo = 0u, hi = 0x3f800000u; // byte(1.0f) = 255 >= b
            while (lo < hi) {
                const uint32_t mid = lo + (hi - lo) / 2;
                float v;
                memcpy(&v, &mid, 4);
                if (tone_byte_host(v) >= b) hi = mid;
                else lo = mid + 1;
            }
            memcpy(&T.t[b], &lo, 4);
        }
        built = true;
    }
    return T;
}
// for tests: the table the kernel searches
extern "C" int ptb_tone_thresholds(float out[256]) {
    if (!out) return PTB_ERR_INVALID;
    memcpy(out, tone_thresholds().t, sizeof(float) * 256);
    return PTB_OK;
}

extern "C" int ptb_read_rgba8(ptb_ctx *c, uint8_t *rgba8, int tone_mapping) {
    GUARD(c);
    SETTLE(c);
    if (!rgba8 || !c->F.image) return fail(c, PTB_ERR_INVALID, "no frame / null buffer");
    const uint32_t n = (uint32_t)c->W * c->H;
    uchar4 *d = nullptr;
    CK(cudaMalloc(&d, (size_t)n * 4));
    k_tonemap<<<grid_for(c), 256, 0, c->stream>>>(c->F.image, d, n, tone_mapping, tone_thresholds());
    c->counters[PTB_CNT_KERNELS]++;
    cudaError_t e = cudaMemcpyAsync(rgba8, d, (size_t)n * 4, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    cudaFree(d);
    CK(e);
    return PTB_OK;
}

extern "C" int ptb_read_counters(ptb_ctx *c, uint64_t out[PTB_CNT__N]) {
    GUARD(c);
    SETTLE(c);
    if (!out) return fail(c, PTB_ERR_INVALID, "null");
    memcpy(out, c->counters, sizeof c->counters);
    return PTB_OK;
}

extern "C" void *ptb_image_device_ptr(ptb_ctx *c) { return c ? (void *)c->F.image : nullptr; }

// =============================================================================================
// Subgrid / Subbeam structures (host builders in ptb_grid_build.cpp)
// =============================================================================================
static void leaf_positions(const ptb_ctx *c, std::vector<float> &out) {
    const size_t n = c->mat.size();
    out.resize(9 * n);
    for (size_t k = 0; k < n; ++k) memcpy(&out[9 * k], &c->pos9[9 * (size_t)c->bvh.order[k]], 36);
}

// A tree built by ptb_build_bvh_gpu exists only in HBM; the Subgrid builder walks it on the host, so mirror the
// nodes once and derive the leaf-order range of every child (ranges are contiguous: children cover
// [first, first+count) of their parent's range, a leaf reference carries its own).
static int mirror_gpu_bvh(ptb_ctx *c) {
    if (!c->bvhOnGpu || c->bvh.nodes.size() == c->bvhNodes) return PTB_OK;
    c->bvh.nodes.resize(c->bvhNodes);
    if (c->bvhNodes) CK(cudaMemcpy(c->bvh.nodes.data(), c->dNodes, c->bvhNodes * sizeof(BvhNode64), cudaMemcpyDeviceToHost));
    c->bvh.ranges.assign(2 * c->bvhNodes, BvhRange{0, 0});
    auto leaf_range = [](int32_t ref) {
        const uint32_t r = (uint32_t)~ref;
        return BvhRange{r >> 4, (r & 15u) + 1u};
    };
    // children have larger indices than their parent only in the SAH builder; here use an explicit post-order
    std::vector<std::pair<int32_t, int>> st;
    if (c->bvh.root >= 0) st.push_back({c->bvh.root, 0});
    while (!st.empty()) {
        auto &[node, phase] = st.back();
        const BvhNode64 &nd = c->bvh.nodes[node];
        if (phase == 0) {
            phase = 1;
            const int32_t a = nd.c0, b = nd.c1;
            if (a >= 0) st.push_back({a, 0});
            if (b >= 0) st.push_back({b, 0});
            continue;
        }
        for (int w = 0; w < 2; ++w) {
            const int32_t ch = w ? nd.c1 : nd.c0;
            BvhRange r;
            if (ch < 0) r = leaf_range(ch);
            else {
                const BvhRange &a = c->bvh.ranges[2 * (size_t)ch], &b = c->bvh.ranges[2 * (size_t)ch + 1];
                r = BvhRange{std::min(a.first, b.first), a.count + b.count};
            }
            c->bvh.ranges[2 * (size_t)node + w] = r;
        }
        st.pop_back();
    }
    return PTB_OK;
}

static int ptb_build_subgrid_impl(ptb_ctx *c, uint32_t gs, uint32_t os, uint32_t m) {
    GUARD(c);
    if (!c->bvhBuilt) return fail(c, PTB_ERR_NOT_BUILT, "ptb_build_bvh first (the Subgrid hierarchy derives from the BVH)");
    if (gs < 1 || gs > 64 || os < 1 || os > 64) return fail(c, PTB_ERR_INVALID, "grid sizes must be 1..64");
    // one grid level per 4 BVH levels below the root grid: the walk's slice stack must hold the deepest chain
    if (1 + (c->bvh.max_depth + 3) / 4 > PTB_GDEPTH) return fail(c, PTB_ERR_INVALID, "BVH too deep for the Subgrid walk's stack (PTB_GDEPTH)");
    if (int r = mirror_gpu_bvh(c)) return r;
    std::vector<float> leaf;
    leaf_positions(c, leaf);
    SubgridResult R;
    build_subgrid(leaf.data(), c->mat.size(), c->bvh, gs, os, m, R);
    if (R.lut.size() >= (1ull << 32) || R.idx.size() >= (1ull << 32)) return fail(c, PTB_ERR_INVALID, "Subgrid pools exceed 2^32 entries");
    devFree(c->dGridHdr); devFree(c->dGridLut); devFree(c->dGridIdx);
    CK(cudaMalloc(&c->dGridHdr, std::max<size_t>(1, R.hdr.size()) * sizeof(GridHdr)));
    CK(cudaMalloc(&c->dGridLut, std::max<size_t>(1, R.lut.size()) * sizeof(GridCell)));
    CK(cudaMalloc(&c->dGridIdx, std::max<size_t>(1, R.idx.size()) * sizeof(uint32_t)));
    CK(cudaMemcpy(c->dGridHdr, R.hdr.data(), R.hdr.size() * sizeof(GridHdr), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(c->dGridLut, R.lut.data(), R.lut.size() * sizeof(GridCell), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(c->dGridIdx, R.idx.data(), R.idx.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
    c->S.grid_hdr = c->dGridHdr;
    c->S.grid_lut = c->dGridLut;
    c->S.grid_idx = c->dGridIdx;
    c->gridBytes = R.hdr.size() * sizeof(GridHdr) + R.lut.size() * sizeof(GridCell) + R.idx.size() * sizeof(uint32_t);
    c->gridCounts[0] = R.hdr.size(); c->gridCounts[1] = R.lut.size(); c->gridCounts[2] = R.idx.size();
    c->gridBuilt = true;
    c->S.ref_grids = 0;
    return PTB_OK;
}

// constructGrid with the channel population on the device (ptb_grid_gpu.cu): the hierarchy is walked on the host
// (microseconds per grid), every (grid, channel) pair is one GPU thread.  Produces the same pools as
// ptb_build_subgrid, byte for byte (tests/test_gpu_builders.py compares the digests).
static int ptb_build_subgrid_gpu_impl(ptb_ctx *c, uint32_t gs, uint32_t os, uint32_t m) {
    GUARD(c);
    SETTLE(c);
    if (!c->bvhBuilt) return fail(c, PTB_ERR_NOT_BUILT, "ptb_build_bvh first (the Subgrid hierarchy derives from the BVH)");
    if (gs < 1 || gs > 64 || os < 1 || os > 64) return fail(c, PTB_ERR_INVALID, "grid sizes must be 1..64");
    // one grid level per 4 BVH levels below the root grid: the walk's slice stack must hold the deepest chain
    if (1 + (c->bvh.max_depth + 3) / 4 > PTB_GDEPTH) return fail(c, PTB_ERR_INVALID, "BVH too deep for the Subgrid walk's stack (PTB_GDEPTH)");
    if (int r = mirror_gpu_bvh(c)) return r;
    const auto t0 = std::chrono::steady_clock::now();
    SubgridTopo T;
    subgrid_topology(c->mat.size(), c->bvh, gs, os, m, T);
    devFree(c->dGridHdr); devFree(c->dGridLut); devFree(c->dGridIdx);
    c->gridBuilt = false;
    SubgridDev D;
    std::string msg;
    const cudaError_t e = subgrid_build_gpu(T, c->dVtx, c->dIdOfTri, c->stream, D, msg);
    if (e != cudaSuccess) {
        if (e != cudaErrorInvalidValue) c->poisoned = true;
        return fail(c, e == cudaErrorInvalidValue ? PTB_ERR_INVALID : PTB_ERR_CUDA, msg);
    }
    c->dGridHdr = D.hdr;
    c->dGridLut = D.lut;
    c->dGridIdx = D.idx;
    c->S.grid_hdr = c->dGridHdr;
    c->S.grid_lut = c->dGridLut;
    c->S.grid_idx = c->dGridIdx;
    c->gridBytes = D.n_grids * sizeof(GridHdr) + D.n_cells * sizeof(GridCell) + D.n_idx * sizeof(uint32_t);
    c->gridCounts[0] = D.n_grids; c->gridCounts[1] = D.n_cells; c->gridCounts[2] = D.n_idx;
    c->gridBuilt = true;
    c->S.ref_grids = 0;
    c->lastBuildMs = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    c->lastBuildDeviceMs = D.ms;
    return PTB_OK;
}

// FNV-1a over the bytes of a structure's device pools, in a fixed order — lets a caller (the tests) check that two
// builders produced the same structure without exposing the layouts.  0 if the structure is not built.
extern "C" uint64_t ptb_structure_digest(ptb_ctx *c, int mode) {
    if (!c || c->poisoned || cudaSetDevice(c->device) != cudaSuccess) return 0;
    if (c->pendingRender && finish_mega(c)) return 0;
    uint64_t h = 1469598103934665603ull;
    std::vector<unsigned char> buf;
    auto eat = [&](const void *dev, size_t bytes) {
        const size_t chunk = 256u << 20;
        for (size_t off = 0; off < bytes; off += chunk) {
            const size_t nb = std::min(chunk, bytes - off);
            buf.resize(nb);
            if (cudaMemcpy(buf.data(), (const char *)dev + off, nb, cudaMemcpyDeviceToHost) != cudaSuccess) return false;
            for (size_t i = 0; i < nb; ++i) {
                h ^= buf[i];
                h *= 1099511628211ull;
            }
        }
        return true;
    };
    const size_t n = c->mat.size();
    if (mode == PTB_MODE_BVH && c->bvhBuilt) {
        if (!eat(c->dNodes, c->bvhNodes * sizeof(BvhNode64)) || !eat(c->dTri, 3 * n * sizeof(float4)) || !eat(c->dNrm, 3 * n * sizeof(float4))) return 0;
        return h;
    }
    if (mode == PTB_MODE_TWO_PLANE && c->gridBuilt) {
        if (!eat(c->dGridHdr, c->gridCounts[0] * sizeof(GridHdr)) || !eat(c->dGridLut, c->gridCounts[1] * sizeof(GridCell)) ||
            !eat(c->dGridIdx, c->gridCounts[2] * sizeof(uint32_t)))
            return 0;
        return h;
    }
    if (mode == PTB_MODE_SUB_BEAMS && c->beamBuilt) {
        if (!eat(c->dBeam, c->beamCounts[0] * sizeof(BeamRec)) || !eat(c->dBeamIdx, c->beamCounts[1] * sizeof(uint32_t))) return 0;
        for (int k = 0; k < 3; ++k)
            if (!eat(c->dBeamRoot[k], c->beamCounts[2] * sizeof(uint32_t))) return 0;
        for (const void *host : {(const void *)c->S.beam_gmin, (const void *)c->S.beam_gmax}) // the planes the records refer to
            for (size_t i = 0; i < sizeof c->S.beam_gmin; ++i) {
                h ^= ((const unsigned char *)host)[i];
                h *= 1099511628211ull;
            }
        return h;
    }
    return 0;
}
extern "C" double ptb_last_build_device_ms(const ptb_ctx *c) { return c ? c->lastBuildDeviceMs : 0.0; }

static int ptb_build_subbeams_impl(ptb_ctx *c, uint32_t gs, uint32_t m) {
    GUARD(c);
    if (!c->bvhBuilt) return fail(c, PTB_ERR_NOT_BUILT, "ptb_build_bvh first (it lays out the triangles)");
    if (gs < 1 || gs > 64) return fail(c, PTB_ERR_INVALID, "grid size must be 1..64");
    std::vector<float> leaf;
    leaf_positions(c, leaf);
    SubbeamResult R;
    build_subbeams(leaf.data(), c->mat.size(), c->bvh.lo, c->bvh.hi, gs, m, c->beamMaxDepth, R);
    if (R.beams.size() >= (1ull << 32) || R.idx.size() >= (1ull << 32))
        return fail(c, PTB_ERR_INVALID, "Subbeam pools exceed 2^32 entries: lower beam_max_depth (ptb_set_build_limits) or the grid size");
    devFree(c->dBeam); devFree(c->dBeamIdx);
    for (int k = 0; k < 3; ++k) devFree(c->dBeamRoot[k]);
    CK(cudaMalloc(&c->dBeam, std::max<size_t>(1, R.beams.size()) * sizeof(BeamRec)));
    CK(cudaMalloc(&c->dBeamIdx, std::max<size_t>(1, R.idx.size()) * sizeof(uint32_t)));
    CK(cudaMemcpy(c->dBeam, R.beams.data(), R.beams.size() * sizeof(BeamRec), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(c->dBeamIdx, R.idx.data(), R.idx.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
    c->beamBytes = R.beams.size() * sizeof(BeamRec) + R.idx.size() * sizeof(uint32_t);
    for (int k = 0; k < 3; ++k) {
        CK(cudaMalloc(&c->dBeamRoot[k], std::max<size_t>(1, R.root[k].size()) * sizeof(uint32_t)));
        CK(cudaMemcpy(c->dBeamRoot[k], R.root[k].data(), R.root[k].size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
        c->S.beam_root[k] = c->dBeamRoot[k];
        c->beamBytes += R.root[k].size() * sizeof(uint32_t);
        for (int j = 0; j < 3; ++j) {
            c->S.beam_gmin[k][j] = R.gmin[k][j];
            c->S.beam_gmax[k][j] = R.gmax[k][j];
        }
        c->S.beam_amin[k] = R.amin[k];
        c->S.beam_amax[k] = R.amax[k];
    }
    c->S.beam = reinterpret_cast<const float4 *>(c->dBeam);
    c->S.beam_idx = c->dBeamIdx;
    c->S.beam_grid_size = R.grid_size;
    c->S.n_beams = (uint32_t)R.beams.size();
    c->beamCounts[0] = R.beams.size(); c->beamCounts[1] = R.idx.size(); c->beamCounts[2] = R.root[0].size();
    c->beamBuilt = true;
    return PTB_OK;
}

// buildGridsWithSubBeams on the device (ptb_beam_gpu.cu): one data-parallel step per tree depth instead of the host's
// depth-first recursion; same pools, byte for byte (ptb_structure_digest, tests/test_gpu_subbeam_build.py).
static int ptb_build_subbeams_gpu_impl(ptb_ctx *c, uint32_t gs, uint32_t m) {
    GUARD(c);
    SETTLE(c);
    if (!c->bvhBuilt) return fail(c, PTB_ERR_NOT_BUILT, "ptb_build_bvh first (it lays out the triangles)");
    if (gs < 1 || gs > 64) return fail(c, PTB_ERR_INVALID, "grid size must be 1..64");
    const auto t0 = std::chrono::steady_clock::now();
    devFree(c->dBeam); devFree(c->dBeamIdx);
    for (int k = 0; k < 3; ++k) devFree(c->dBeamRoot[k]);
    c->beamBuilt = false;
    SubbeamDev D;
    std::string msg;
    const cudaError_t e = subbeams_build_gpu(c->dVtx, c->dIdOfTri, c->mat.size(), c->bvh.lo, c->bvh.hi, gs, m, c->beamMaxDepth, c->stream, D, msg);
    if (e != cudaSuccess) {
        if (e != cudaErrorInvalidValue) c->poisoned = true;
        return fail(c, e == cudaErrorInvalidValue ? PTB_ERR_INVALID : PTB_ERR_CUDA,
                    e == cudaErrorInvalidValue ? msg + ": lower beam_max_depth (ptb_set_build_limits) or the grid size" : msg);
    }
    c->dBeam = D.beams;
    c->dBeamIdx = D.idx;
    const size_t cells = (size_t)gs * gs * gs * gs;
    std::vector<uint32_t> ids(cells);
    c->beamBytes = D.n_beams * sizeof(BeamRec) + D.n_idx * sizeof(uint32_t) + 3 * cells * sizeof(uint32_t);
    for (int k = 0; k < 3; ++k) {
        for (size_t i = 0; i < cells; ++i) ids[i] = (uint32_t)(k * cells + i); // root beams live at [axis * cells + cell]
        CK(cudaMalloc(&c->dBeamRoot[k], cells * sizeof(uint32_t)));
        CK(cudaMemcpy(c->dBeamRoot[k], ids.data(), cells * sizeof(uint32_t), cudaMemcpyHostToDevice));
        c->S.beam_root[k] = c->dBeamRoot[k];
        for (int j = 0; j < 3; ++j) {
            c->S.beam_gmin[k][j] = D.gmin[k][j];
            c->S.beam_gmax[k][j] = D.gmax[k][j];
        }
        c->S.beam_amin[k] = D.amin[k];
        c->S.beam_amax[k] = D.amax[k];
    }
    c->S.beam = reinterpret_cast<const float4 *>(c->dBeam);
    c->S.beam_idx = c->dBeamIdx;
    c->S.beam_grid_size = gs;
    c->S.n_beams = (uint32_t)D.n_beams;
    c->beamCounts[0] = D.n_beams; c->beamCounts[1] = D.n_idx; c->beamCounts[2] = cells;
    c->beamBuilt = true;
    c->lastBuildMs = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    c->lastBuildDeviceMs = D.ms;
    return PTB_OK;
}

// =============================================================================================
// ptb_upload_subgrid: the reference's own Subgrid hierarchy
// =============================================================================================
// Takes std::vector<Grid> (lightFieldGrid.h:10-22) flattened — one ptb_ref_grid per Grid plus its three vectors concatenated
// in grid order — and lays it out in the device pools subgrid_closest walks (64-byte header, {start,end} cells, index pool).
// intersectGrid (lightFieldGrid.cpp:71-147) and subgrid_closest are the same statements, so over these grids the two test
// counters and the heat maps are the reference's numbers.  The input is untrusted: every range, link and the nesting depth
// are checked before anything reaches the device.
static int ptb_upload_subgrid_impl(ptb_ctx *c, const ptb_ref_grid *grids, uint32_t n_grids, const uint32_t *lut_start, const uint32_t *lut_end,
                                   uint64_t n_cells, const uint32_t *indices, uint64_t n_indices) {
    GUARD(c);
    SETTLE(c);
    if (!c->bvhBuilt) return fail(c, PTB_ERR_NOT_BUILT, "ptb_build_bvh first (it lays out the triangles the grids index)");
    if (!grids || n_grids < 3) return fail(c, PTB_ERR_INVALID, "reference grids: at least the three axis roots are needed");
    if ((n_cells && (!lut_start || !lut_end)) || (n_indices && !indices)) return fail(c, PTB_ERR_INVALID, "reference grids: null pools");
    if (n_cells >= (1ull << 32) || n_indices >= (1ull << 32)) return fail(c, PTB_ERR_INVALID, "reference grids: pools exceed 2^32 entries");
    const uint64_t nTris = c->mat.size();
    std::vector<GridHdr> hdr(n_grids);
    std::vector<GridCell> lut((size_t)n_cells);
    std::vector<uint32_t> idx((size_t)n_indices);
    std::vector<int> depth(n_grids, 1); // inner grids nested below and including g (the slice stack of subgrid_closest_slices)
    uint64_t cellAt = 0, idxAt = 0;
    std::vector<uint64_t> cellBase(n_grids), idxBase(n_grids);
    for (uint32_t g = 0; g < n_grids; ++g) {
        const ptb_ref_grid &G = grids[g];
        const std::string at = "reference grid " + std::to_string(g) + ": ";
        if (G.size < 1 || G.size > 256) return fail(c, PTB_ERR_INVALID, at + "size out of range");
        const uint64_t want = (uint64_t)G.size * G.size * G.size * G.size;
        if (G.n_cells != want) return fail(c, PTB_ERR_INVALID, at + "LUT has " + std::to_string(G.n_cells) + " cells, size^4 = " + std::to_string(want));
        if (G.splitting_axis < 0 || G.splitting_axis > 2) return fail(c, PTB_ERR_INVALID, at + "splitting axis");
        if (cellAt + G.n_cells > n_cells || idxAt + G.n_indices > n_indices) return fail(c, PTB_ERR_INVALID, at + "pools shorter than the grids claim");
        // the walk derives grid.max[axis] as aabb.max[axis] + 0.1f (adjustGridSize, lightFieldGrid.cpp:176-178): hold the input to it
        const float derived = G.aabb_max[G.splitting_axis] + 0.1f;
        if (memcmp(&derived, &G.max[G.splitting_axis], 4) != 0)
            return fail(c, PTB_ERR_INVALID, at + "max[axis] is not aabb.max[axis] + 0.1f (not a grid constructGrid built)");
        cellBase[g] = cellAt;
        idxBase[g] = idxAt;
        GridHdr &H = hdr[g];
        for (int k = 0; k < 3; ++k) {
            H.amin[k] = G.aabb_min[k];
            H.amax[k] = G.aabb_max[k];
            H.gmin[k] = G.min[k];
            H.inv_delta[k] = G.inv_delta[k];
        }
        H.size_flags = G.size | (G.has_tris ? 0x80000000u : 0u);
        H.lut_base = (uint32_t)cellAt;
        H.pad0 = H.pad1 = 0;
        for (uint64_t k = 0; k < G.n_cells; ++k) {
            const uint32_t s0 = lut_start[cellAt + k], e0 = lut_end[cellAt + k];
            if (s0 > e0 || e0 > G.n_indices) return fail(c, PTB_ERR_INVALID, at + "LUT cell outside the grid's index list");
            lut[cellAt + k] = GridCell{(uint32_t)(idxAt + s0), (uint32_t)(idxAt + e0)};
        }
        for (uint64_t k = 0; k < G.n_indices; ++k) {
            const uint32_t v = indices[idxAt + k];
            if (G.has_tris) {
                if (v >= nTris) return fail(c, PTB_ERR_INVALID, at + "triangle id out of range");
                idx[idxAt + k] = c->triOfId[v]; // reference triangle id -> leaf position
            } else {
                // children are constructed after their parent (lightFieldGrid.cpp:857-859), object grids after the roots: a
                // larger id — which also rules out cycles
                if (v >= n_grids || v <= g) return fail(c, PTB_ERR_INVALID, at + "child grid id out of range or not after its parent");
                if (grids[v].splitting_axis != G.splitting_axis) return fail(c, PTB_ERR_INVALID, at + "child grid of another axis");
                idx[idxAt + k] = v;
            }
        }
        cellAt += G.n_cells;
        idxAt += G.n_indices;
    }
    if (cellAt != n_cells || idxAt != n_indices) return fail(c, PTB_ERR_INVALID, "reference grids: pools longer than the grids claim");
    for (uint32_t g = n_grids; g-- > 0;) { // children have larger ids: their depth is final when the parent is reached
        if (grids[g].has_tris) continue;
        int d = 0;
        for (uint64_t k = 0; k < grids[g].n_indices; ++k) d = std::max(d, depth[indices[idxBase[g] + k]]);
        depth[g] = d + 1;
    }
    const int deepest = std::max(depth[0], std::max(depth[1], depth[2]));
    if (deepest > PTB_GDEPTH)
        return fail(c, PTB_ERR_INVALID, "reference grids nest " + std::to_string(deepest) + " deep, the walk's stack holds " + std::to_string(PTB_GDEPTH));
    devFree(c->dGridHdr); devFree(c->dGridLut); devFree(c->dGridIdx);
    c->gridBuilt = false;
    CK(cudaMalloc(&c->dGridHdr, hdr.size() * sizeof(GridHdr)));
    CK(cudaMalloc(&c->dGridLut, std::max<size_t>(1, lut.size()) * sizeof(GridCell)));
    CK(cudaMalloc(&c->dGridIdx, std::max<size_t>(1, idx.size()) * sizeof(uint32_t)));
    CK(cudaMemcpy(c->dGridHdr, hdr.data(), hdr.size() * sizeof(GridHdr), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(c->dGridLut, lut.data(), lut.size() * sizeof(GridCell), cudaMemcpyHostToDevice));
    CK(cudaMemcpy(c->dGridIdx, idx.data(), idx.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
    c->S.grid_hdr = reinterpret_cast<const float4 *>(c->dGridHdr);
    c->S.grid_lut = c->dGridLut;
    c->S.grid_idx = c->dGridIdx;
    c->gridBytes = hdr.size() * sizeof(GridHdr) + lut.size() * sizeof(GridCell) + idx.size() * sizeof(uint32_t);
    c->gridCounts[0] = hdr.size(); c->gridCounts[1] = lut.size(); c->gridCounts[2] = idx.size();
    // getMemory2Plane (lightFieldGrid.cpp:60-68) of what was handed over: 288-byte Grid objects + three u32 vectors
    c->refGridBytes = (uint64_t)n_grids * 288u + 4u * (2u * n_cells + n_indices);
    c->gridBuilt = true;
    c->S.ref_grids = 1; // re-traced misses add their (identical) tests again, like the CPU does (tracer.cpp:127-132)
    return PTB_OK;
}
extern "C" uint64_t ptb_ref_subgrid_bytes(const ptb_ctx *c) { return (c && c->gridBuilt && c->S.ref_grids) ? c->refGridBytes : 0; }

// =============================================================================================
// exception barrier of the entry points that allocate from caller- or file-supplied sizes
// =============================================================================================
// C++ exceptions (std::bad_alloc / length_error from a corrupt count, anything a builder throws) must not cross the extern "C"
// boundary: under ctypes or inside the facade they would abort the host process.  They become a status + ptb_last_error.
template <typename F>
static int guarded(ptb_ctx *c, F &&f) {
    try {
        return f();
    } catch (const std::bad_alloc &) {
        return fail(c, PTB_ERR_INVALID, "out of host memory (corrupt element count?)");
    } catch (const std::exception &e) {
        return fail(c, PTB_ERR_INVALID, std::string("exception: ") + e.what());
    } catch (...) {
        return fail(c, PTB_ERR_INVALID, "unknown C++ exception");
    }
}
extern "C" int ptb_upload_triangles(ptb_ctx *c, const float *pos9, const float *nrm9, const float *uv6, const int32_t *mat, uint64_t n) {
    return guarded(c, [&]() { return ptb_upload_triangles_impl(c, pos9, nrm9, uv6, mat, n); });
}
extern "C" int ptb_upload_objects(ptb_ctx *c, const ptbs_object *objects, uint32_t n) {
    return guarded(c, [&]() { return ptb_upload_objects_impl(c, objects, n); });
}
extern "C" int ptb_upload_materials(ptb_ctx *c, const ptbs_material *materials, uint32_t n) {
    return guarded(c, [&]() { return ptb_upload_materials_impl(c, materials, n); });
}
extern "C" int ptb_upload_textures(ptb_ctx *c, const ptb_texture *textures, uint32_t n_textures, const int32_t *material_tex2, uint32_t n_materials) {
    return guarded(c, [&]() { return ptb_upload_textures_impl(c, textures, n_textures, material_tex2, n_materials); });
}
extern "C" int ptb_load_ptbs(ptb_ctx *c, const char *path) {
    return guarded(c, [&]() { return ptb_load_ptbs_impl(c, path); });
}
extern "C" int ptb_load_scene_file(ptb_ctx *c, const char *scene_file, const char *base_dir, int flags) {
    return guarded(c, [&]() { return ptb_load_scene_file_impl(c, scene_file, base_dir, flags); });
}
extern "C" int ptb_build_bvh(ptb_ctx *c) {
    return guarded(c, [&]() { return ptb_build_bvh_impl(c); });
}
extern "C" int ptb_build_bvh_gpu(ptb_ctx *c) {
    return guarded(c, [&]() { return ptb_build_bvh_gpu_impl(c); });
}
extern "C" int ptb_upload_ref_bvh(ptb_ctx *c, const ptb_ref_bvh_node *nodes, uint32_t n_nodes, const uint32_t *indices, uint32_t n_indices,
                                  const int32_t *object_roots, uint32_t n_objects, int32_t scene_root) {
    return guarded(c, [&]() { return ptb_upload_ref_bvh_impl(c, nodes, n_nodes, indices, n_indices, object_roots, n_objects, scene_root); });
}
extern "C" int ptb_build_subgrid(ptb_ctx *c, uint32_t gs, uint32_t os, uint32_t m) {
    return guarded(c, [&]() { return ptb_build_subgrid_impl(c, gs, os, m); });
}
extern "C" int ptb_build_subgrid_gpu(ptb_ctx *c, uint32_t gs, uint32_t os, uint32_t m) {
    return guarded(c, [&]() { return ptb_build_subgrid_gpu_impl(c, gs, os, m); });
}
extern "C" int ptb_build_subbeams(ptb_ctx *c, uint32_t gs, uint32_t m) {
    return guarded(c, [&]() { return ptb_build_subbeams_impl(c, gs, m); });
}
extern "C" int ptb_build_subbeams_gpu(ptb_ctx *c, uint32_t gs, uint32_t m) {
    return guarded(c, [&]() { return ptb_build_subbeams_gpu_impl(c, gs, m); });
}

// =============================================================================================
// ptb_trace_rays_static: ptb_trace_rays without any allocation
// =============================================================================================
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  ptb_trace_rays stages the host rays and
// the results in per-context scratch that ensure_scratch cudaMalloc's on demand; here the same kernel (k_trace_bvh: the
// persistent dynamic-fetch BVH walk) runs on scratch that lives in ONE file-scope device array of fixed size:
//
//   [0, 256)                      the launch's work cursor (8 bytes used; must start at zero, cleared per call)
//   + 24 B x PTB_STATIC_MAX_RAYS  staged rays (origin, direction)
//   +  4 B x PTB_STATIC_MAX_RAYS  hit distances
//   +  4 B x PTB_STATIC_MAX_RAYS  hit triangle ids
//
// = 2 MiB + 256 B at PTB_STATIC_MAX_RAYS = 65,536, held by every GPU that loads the library until the process exits.
// All calls — from any context on that device — share the array, and it keeps the last call's contents: the call clears
// the cursor and overwrites [0, n) of the other regions before the kernel reads them, and it is synchronous, so calls
// issued one after the other are safe; calls that overlap on the GPU (two host threads, two contexts of one device) race.
//
// In host code `g_staticScratch` is a host-side stand-in for the symbol, NOT a device address: its device address is
// resolved once per device with cudaGetSymbolAddress (that device current) and only that pointer is handed to copies
// and kernels.
static __device__ __align__(256) unsigned char g_staticScratch[256 + (size_t)PTB_STATIC_MAX_RAYS * 32];

namespace {
const int kMaxDevices = 64;
unsigned char *g_staticScratchPtr[kMaxDevices] = {nullptr}; // device address of g_staticScratch on device i (0 = not resolved yet)
} // namespace

extern "C" uint64_t ptb_trace_rays_static_max(void) { return (uint64_t)PTB_STATIC_MAX_RAYS; }

extern "C" int ptb_trace_rays_static(ptb_ctx *c, const float *od6, uint64_t n, float *t, int32_t *tri) {
    GUARD(c); // makes c->device current
    SETTLE(c);
    if (n > (uint64_t)PTB_STATIC_MAX_RAYS)
        return fail(c, PTB_ERR_INVALID, "ptb_trace_rays_static: " + std::to_string(n) + " rays exceed the static scratch of " +
                                            std::to_string(PTB_STATIC_MAX_RAYS) + " rays (PTB_STATIC_MAX_RAYS); split the batch or use ptb_trace_rays");
    int r = check_mode_ready(c, PTB_MODE_BVH);
    if (r) return r;
    if (n == 0) return PTB_OK;
    if (!od6 || !t || !tri) return fail(c, PTB_ERR_INVALID, "null ray / result pointers");
    if (c->device < 0 || c->device >= kMaxDevices) return fail(c, PTB_ERR_INVALID, "device index beyond the static-scratch table");
    unsigned char *base = g_staticScratchPtr[c->device];
    if (!base) { // set-up, once per device, with that device current
        void *p = nullptr;
        CK(cudaGetSymbolAddress(&p, g_staticScratch));
        if (!p) return fail(c, PTB_ERR_CUDA, "cudaGetSymbolAddress returned a null address for the static scratch");
        g_staticScratchPtr[c->device] = base = (unsigned char *)p;
    }
    unsigned long long *cursor = reinterpret_cast<unsigned long long *>(base);
    float *dRays = reinterpret_cast<float *>(base + 256);
    float *dT = reinterpret_cast<float *>(base + 256 + (size_t)PTB_STATIC_MAX_RAYS * 24);
    int *dTri = reinterpret_cast<int *>(base + 256 + (size_t)PTB_STATIC_MAX_RAYS * 28);
    CK(cudaMemsetAsync(cursor, 0, 256, c->stream)); // what must start at zero
    CK(cudaMemcpyAsync(dRays, od6, n * 6 * sizeof(float), cudaMemcpyHostToDevice, c->stream));
    const int gp = (int)std::min<uint64_t>((uint64_t)c->sms * c->extendCtas, (n + 127) / 128);
    CK(cudaEventRecord(c->evStart, c->stream));
    k_trace_bvh<false, false><<<gp, 128, 0, c->stream>>>(c->S, dRays, n, dT, dTri, c->dIdOfTri, c->dCounters, cursor, c->fetchKeep, 0);
    CK(cudaGetLastError());
    CK(cudaEventRecord(c->evStop, c->stream));
    CK(cudaMemcpyAsync(t, dT, n * sizeof(float), cudaMemcpyDeviceToHost, c->stream));
    CK(cudaMemcpyAsync(tri, dTri, n * sizeof(int), cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    float ms = 0;
    cudaEventElapsedTime(&ms, c->evStart, c->evStop);
    c->lastMs = ms;
    c->counters[PTB_CNT_KERNELS]++;
    c->counters[PTB_CNT_RAYS] += n;
    return PTB_OK;
}
extern "C" int ptb_upload_subgrid(ptb_ctx *c, const ptb_ref_grid *grids, uint32_t n_grids, const uint32_t *lut_start, const uint32_t *lut_end,
                                  uint64_t n_cells, const uint32_t *indices, uint64_t n_indices) {
    return guarded(c, [&]() { return ptb_upload_subgrid_impl(c, grids, n_grids, lut_start, lut_end, n_cells, indices, n_indices); });
}
