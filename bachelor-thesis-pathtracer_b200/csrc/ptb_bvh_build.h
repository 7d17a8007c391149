// ptb_bvh_build.h — host-side construction of the flattened BVH the sm_100a traversal reads.
//
// Stands in for constructBVH (reference: src/accelerationStructures/bvh.cpp:222-390, called from
// src/scene/scene.cpp:113-120).  The reference builds a two-level tree (objects, then triangles
// per object) with a 12-plane x 3-axis sweep that re-bins every primitive per candidate
// (O(36 n) per node); here ONE tree over all triangles is built with a 16-bin SAH sweep
// (O(n) per node), multi-threaded over subtrees, and written directly in the 64-byte two-box
// node layout of DESIGN.md.  Closest hits do not depend on the tree, only visit order does.
#pragma once
#include <cstdint>
#include <vector>

namespace ptb {

struct BvhNode64 { // one inner node = both children's boxes + links; 4 x float4
    float lo0[3], hi0[3], lo1[3], hi1[3];
    int32_t c0, c1; // >= 0: inner node index; < 0: leaf ~((first << 4) | (count - 1))
    int32_t pad0, pad1;
};
static_assert(sizeof(BvhNode64) == 64, "node must be 64 bytes");

struct BvhRange { // leaf-order triangle range [first, first+count) covered by a subtree
    uint32_t first, count;
};

struct BvhResult {
    std::vector<BvhNode64> nodes;
    std::vector<BvhRange> ranges; // 2 per node: ranges of child 0 and child 1 (Subgrid builder input)
    std::vector<uint32_t> order; // leaf position -> reference triangle id
    int32_t root = 0;            // node reference of the root (may be a leaf for tiny scenes)
    int max_depth = 0;
    float lo[3], hi[3];          // scene bounds
};

// pos9: n x (v0,v1,v2).  leaf_max <= 16.
void build_bvh(const float *pos9, uint64_t n, int leaf_max, BvhResult &out);


} // namespace ptb
