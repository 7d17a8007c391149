// structures_facade.cpp — drop-in replacement for the reference's three structure files
//   src/accelerationStructures/bvh.cpp, lightFieldGrid.cpp, lightFieldGridSubBeams.cpp
// behind their own headers (bvh.h:26-38, lightFieldGrid.h:23-27, lightFieldGridSubBeams.h:27-30), for use together with
// tracer_facade.cpp.  Compile it INSTEAD of those three files.  Then
//   * initScene (scene.cpp:113-125) builds on the GPU: its constructBVH / constructGrid calls land here and forward to
//     ptb_build_bvh / ptb_build_subgrid_gpu (the reference needs 26 s + 1001 s for the Lucy stand-in, this needs < 1 s);
//   * main.cpp needs NO added line: setGridSettings / setGridBeamsSettings (main.cpp:91-92) are the functions below;
//   * getMemoryBVH / getMemory2Plane / getMemoryGridBeams (main.cpp:136, window.cpp:200-202) report the structure that is
//     actually traversed (device bytes).
// The CPU traversal entry points (findBVHIntesection, intersectGrid, intersectSubBeamGrid) stay callable: nothing on the render
// path uses them any more (trace() runs on the GPU), but scene.cpp's findIntersection links against them, so a stray caller
// gets a correct answer — the closest hit found on the GPU (ptb_trace_rays, one ray) and then the reference's OWN
// triangleIntersection (triangle.cpp:7-52) on that one triangle, which fills tmax / normal / tangent / uv / material exactly as
// the CPU walk would have.  The two per-ray test counters are not advanced by such calls.
// What does not exist in this flavour: the reference's node array.  getNode / getNodeAABB / bvhGetTrisIndex return an empty
// leaf, and ptbFacadeUseReferenceTree (count-exact traversal of the reference's tree) needs the reference's bvh.cpp linked.
#include <cstdio>

#include "accelerationStructures/bvh.h"
#include "accelerationStructures/lightFieldGrid.h"
#include "accelerationStructures/lightFieldGridSubBeams.h"
#include "common.h"
#include "primitives/triangle.h"
#include "scene/scene.h"

#include "facade_internal.h"
#include "tracer_facade.h"

namespace {
BvhSettings settings{.maxDepth = 0};
BvhNode emptyNode{.startIdx = 0, .endIdx = 0, .splitAxis = 0, .depth = 0, .AABBIdx = 0, .childLeft = -1, .childRight = -1, .cost = 0.0f, .hasTris = true};
AABB emptyBox{};
u32 zeroIndex = 0;

void noTree(const char *what) {
    static bool said = false;
    if (!said) fprintf(stderr, "ptb facade: %s: the reference's node array does not exist when structures_facade.cpp fronts bvh.h\n", what);
    said = true;
}

// closest hit of ONE ray on the GPU, attributes by the reference's own triangle test
bool traceOne(Ray &ray, int mode) {
    if (!ptbfacade::structureOnDevice(mode)) return false;
    const float od[6] = {ray.origin[0], ray.origin[1], ray.origin[2], ray.direction[0], ray.direction[1], ray.direction[2]};
    float t = 0.0f;
    int32_t tri = -1;
    if (ptb_trace_rays(ptbfacade::context(), od, 1, mode, &t, &tri) != 0 || tri < 0 || (size_t)tri >= getTris().size()) return false;
    return triangleIntersection(ray, getTris()[tri]); // accepts only inside the ray's own [tmin, tmax], like the CPU walk
}

struct Announce {
    Announce() { ptbfacade::structuresAreFronted(); }
} announce;
} // namespace

// ---- bvh.h:26-38 -------------------------------------------------------------------------------------------------------
BvhSettings *getBvhSettings() { return &settings; }
BvhNode &getNode(u32) {
    noTree("getNode");
    return emptyNode;
}
AABB &getNodeAABB(u32) {
    noTree("getNodeAABB");
    return emptyBox;
}
u32 &bvhGetTrisIndex(u32) {
    noTree("bvhGetTrisIndex");
    return zeroIndex;
}
void destroyBVH() {}
void calculateBoundingBox(BvhNode &, bool) {}
// scene.cpp:113-118 asks for a scene-level tree over the objects and then one tree per object.  ONE tree over all triangles
// serves every ray here (closest hits do not depend on the tree), built when the scene-level call arrives — the triangles are
// loaded by then (scene.cpp:97-100).  The returned "root" is what findBVHIntesection below gets back: unused.
int constructBVH(int, int, int, bool isObject) {
    if (!isObject) ptbfacade::sceneOnDevice();
    return 0;
}
bool findBVHIntesection(Ray &ray, int) {
    if (ray.terminated) return false; // bvh.cpp:51
    return traceOne(ray, BVH);
}
unsigned long getMemoryBVH() { return ptbFacadeStructureBytes(BVH); }
int getBVHDepth(int) { return ptbfacade::context() ? ptb_bvh_depth(ptbfacade::context()) : 0; }

// ---- lightFieldGrid.h:23-27 --------------------------------------------------------------------------------------------
void intersectGrid(Ray &ray) { traceOne(ray, TWO_PLANE); } // no `terminated` check, like lightFieldGrid.cpp:71
void constructGrid() { ptbfacade::structureOnDevice(TWO_PLANE); }
void constructChannel(float, float, float, float, int, std::vector<u32>) {} // a step of the reference's own builder; nothing to do
void setGridSettings(u32 size, u32 oSize, u32 count) { ptbFacadeSetGridSettings(size, oSize, count); }
unsigned long getMemory2Plane() { return ptbFacadeStructureBytes(TWO_PLANE); }

// ---- lightFieldGridSubBeams.h:27-30 ------------------------------------------------------------------------------------
u64 getMemoryGridBeams() { return ptbFacadeStructureBytes(SUB_BEAMS); }
void setGridBeamsSettings(u32 size, u32 count) { ptbFacadeSetGridBeamsSettings(size, count); }
void buildGridsWithSubBeams() { ptbfacade::structureOnDevice(SUB_BEAMS); }
void intersectSubBeamGrid(Ray &ray) {
    if (ray.terminated) return; // lightFieldGridSubBeams.cpp:131
    traceOne(ray, SUB_BEAMS);
}
