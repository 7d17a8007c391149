// ptb_lbvh.h — construction of the flattened BVH ON THE GPU.
//
// Stands in for constructBVH (reference: src/accelerationStructures/bvh.cpp:116-189,222-390), like
// ptb_bvh_build.h, but for scenes the host builders cannot reach: the reference needs 7.2 s for 1e5
// triangles and grows superlinearly; the host SAH builder of this library keeps a
// 36-byte host copy plus per-subtree vectors per triangle.  Here the whole build — scene bounds, 63-bit
// Morton keys, radix sort, Karras' radix-tree hierarchy, bottom-up boxes, collapse of small subtrees into
// leaves, the 64-byte two-box nodes and the leaf-ordered triangle / attribute arrays the traversal kernels
// read — runs in HBM.  Closest hits do not depend on the tree (ptb_device.cuh resolves equal distances
// by triangle id), so frames are identical to those over the SAH tree; only the visit counts differ.
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <string>

namespace ptb {

struct LbvhResult {
    // device arrays, cudaMalloc'ed by lbvh_build; the caller owns (cudaFree) them
    float4 *tri = nullptr;        // 3 x float4 per triangle in leaf order: (v0, reference id), (e1,-), (e2,-)
    float4 *nrm = nullptr;        // 3 x float4: (n0, material id), (n1,-), (n2,-)
    float4 *nodes = nullptr;      // n_nodes x 64 B (BvhNode64)
    uint32_t *tri_of_id = nullptr; // reference id -> leaf position
    uint32_t *id_of_tri = nullptr; // leaf position -> reference id
    uint64_t n_nodes = 0;
    int32_t root = 0;  // node reference of the root (a leaf reference when n <= leaf_max)
    int height = 0;    // inner nodes on the longest root-to-leaf chain (bounds the traversal stack)
    float lo[3] = {0, 0, 0}, hi[3] = {0, 0, 0}; // scene bounds
    double ms = 0.0;   // device time of the build (CUDA events on `stream`)
};

// d_vtx: n x 9 floats (v0,v1,v2) by reference id; d_nrm9 likewise; d_mat: n material indices — all device
// pointers.  1 <= leaf_max <= 16, n < 2^27.  A tree of height >= max_height (the traversal stack) is rebuilt
// from coarser Morton keys.  Everything is ordered on `stream`; returns after the stream has been synchronised.
// On failure returns the CUDA error (or cudaErrorInvalidValue) and fills `err`; nothing stays allocated.
cudaError_t lbvh_build(const float *d_vtx, const float *d_nrm9, const int32_t *d_mat, uint64_t n, int leaf_max, int max_height,
                       cudaStream_t stream, LbvhResult &out, std::string &err);

} // namespace ptb
