// ptb_grid_gpu.h — construction of the Subgrid ("2Plane"/Hybrid) structure with the channel population ON THE
// GPU.  Stands in for constructGrid / constructChannel
// (reference: src/accelerationStructures/lightFieldGrid.cpp:189-322,381-414,493-583).
//
// The reference fills size^4 channels per grid one after the other, copying its candidate vector per channel
// (64 s for 1e5 triangles); the host builder of this library (ptb_grid_build.cpp) spreads the
// grids of a level over the host threads (30 s for the Lucy stand-in).  The channels are independent: here the
// hierarchy (a cheap walk over the BVH) stays on the host and every (grid, channel) pair becomes one GPU thread —
//   pass 1  separating-axis tests of the channel against the grid's candidates -> count (+ 64-bit accept mask)
//   scan    exclusive sum of the counts = the channel's slice of the index pool
//   pass 2  the accepted candidates, in the reference's order, into the pool; {start,end} into the LUT
// — with the same fp32 expressions as the host builder (ptb_sat.h), so both produce identical pools.
#pragma once
#include <cstdint>
#include <string>
#include <vector>

#include "ptb_bvh_build.h"
#include "ptb_grid_build.h"
#include "ptb_sat.h"

namespace ptb {

struct GridTopo { // one grid of the hierarchy before its channels are filled; 64 bytes
    float lo[3];
    uint32_t axis_size_flags; // axis | size << 8 | hasTris << 31
    float hi[3];
    uint32_t cand_base;  // hasTris: first triangle (leaf order); else: offset of the child ids in the candidate pool
    float gmin[3];       // the enlarged planes (adjustGridSize)
    uint32_t cand_count;
    float gmax[3];
    uint32_t pad;
};
static_assert(sizeof(GridTopo) == 64, "GridTopo must be 64 bytes");

struct SubgridTopo {
    std::vector<GridTopo> grids;     // grids 0..2 are the scene roots for axis 0..2; same numbering as build_subgrid
    std::vector<uint64_t> cell_base; // grids.size() + 1 prefix sums of size^4
    std::vector<uint32_t> cand;      // child grid ids of the inner grids, ordered by aabb.min[axis] (lightFieldGrid.cpp:395-400)
    uint32_t max_cand = 0;
};

// The hierarchy of build_subgrid without the channel population (4 BVH levels per grid level, leaf grids of
// <= m triangles; lightFieldGrid.cpp:216-294).  Implemented in ptb_grid_build.cpp.
void subgrid_topology(uint64_t n, const BvhResult &bvh, uint32_t gs, uint32_t os, uint32_t m, SubgridTopo &out);
// The 64-byte device header of grid g (what build_subgrid's flatten step writes).
GridHdr subgrid_header(const GridTopo &g, uint64_t lut_base);

// One channel: the candidates of `g` that the sheared cuboid of cell `local` touches.  Returns their number;
// sets bit k of *mask for accepted candidate k < 64 (if mask), writes their ids in candidate order (if out).
// tri9 + 9 * (remap ? remap[i] : i) are the vertices of leaf-order triangle i.
PTB_HD uint32_t populate_cell(const GridTopo &g, uint32_t local, const GridTopo *grids, const uint32_t *cand, const float *tri9,
                              const uint32_t *remap, unsigned long long *mask, uint32_t *out) {
    const int axis = (int)(g.axis_size_flags & 0xffu);
    const uint32_t sz = (g.axis_size_flags >> 8) & 0xffffu;
    const bool hasTris = (g.axis_size_flags >> 31) != 0u;
    const uint32_t t = local % sz, s = (local / sz) % sz, v = (local / (sz * sz)) % sz, u = local / (sz * sz * sz);
    const float step = 1.0f / (float)sz;
    const float fu = (float)u / (float)sz, fv = (float)v / (float)sz, fs = (float)s / (float)sz, ft = (float)t / (float)sz;
    Channel ch;
    make_channel(V3{g.gmin[0], g.gmin[1], g.gmin[2]}, V3{g.gmax[0], g.gmax[1], g.gmax[2]}, axis, fu, fv, fu + step, fv + step, fs, ft,
                 fs + step, ft + step, ch);
    uint32_t cnt = 0;
    unsigned long long bits = 0ull;
    for (uint32_t k = 0; k < g.cand_count; ++k) {
        bool in;
        uint32_t id;
        if (hasTris) {
            id = g.cand_base + k;
            const float *p = tri9 + 9 * (size_t)(remap ? remap[id] : id);
            float t9[9];
            PTB_UNROLL
            for (int j = 0; j < 9; ++j) t9[j] = p[j];
            in = tri_in_channel(t9, ch);
        } else {
            id = cand[g.cand_base + k];
            const GridTopo &c = grids[id];
            in = box_in_channel(V3{c.lo[0], c.lo[1], c.lo[2]}, V3{c.hi[0], c.hi[1], c.hi[2]}, ch);
        }
        if (in) {
            if (out) out[cnt] = id;
            if (k < 64u) bits |= 1ull << k;
            cnt++;
        }
    }
    if (mask) *mask = bits;
    return cnt;
}

#ifdef __CUDACC__
struct SubgridDev { // device pools, cudaMalloc'ed by subgrid_build_gpu; the caller owns (cudaFree) them
    float4 *hdr = nullptr;  // 64 B per grid (GridHdr)
    uint2 *lut = nullptr;   // {start,end} per cell
    uint32_t *idx = nullptr;
    uint64_t n_grids = 0, n_cells = 0, n_idx = 0;
    double ms = 0.0;        // device time of the population (CUDA events on `stream`)
};
// d_vtx: 9 floats per triangle by reference id, d_id_of_tri: leaf position -> reference id (device pointers).
cudaError_t subgrid_build_gpu(const SubgridTopo &topo, const float *d_vtx, const uint32_t *d_id_of_tri, cudaStream_t stream,
                              SubgridDev &out, std::string &err);
#endif

} // namespace ptb
