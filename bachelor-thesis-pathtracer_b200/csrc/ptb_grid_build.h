// ptb_grid_build.h — host-side construction of the thesis's light-field structures in the
// flattened device layouts (DESIGN.md "Data layout"):
//   Subgrid / "2Plane" / Hybrid hierarchy  (reference: src/accelerationStructures/lightFieldGrid.cpp:189-583)
//   Subbeams                                (reference: src/accelerationStructures/lightFieldGridSubBeams.cpp:185-374)
// Channel population uses the same separating-axis tests as the reference (src/types/sat.cpp).
#pragma once
#include <cstdint>
#include <vector>

#include "ptb_bvh_build.h"

namespace ptb {

struct GridHdr { // 64 bytes = 4 x float4, read by subgrid_closest()
    float amin[3];
    uint32_t size_flags; // size | hasTris << 31
    float amax[3];
    uint32_t lut_base;   // first cell of this grid in the LUT pool
    float gmin[3];
    uint32_t pad0;
    float inv_delta[3];
    uint32_t pad1;
};
static_assert(sizeof(GridHdr) == 64, "grid header must be 64 bytes");

struct GridCell {
    uint32_t start, end; // absolute range in the index pool
};

struct SubgridResult {
    std::vector<GridHdr> hdr;   // grids 0..2 are the scene roots for axis 0..2
    std::vector<GridCell> lut;
    std::vector<uint32_t> idx;  // child grid ids, or leaf-order triangle indices
};

// tri_leaf: 9 floats per triangle in LEAF order (v0,v1,v2).  gs/os/m as setGridSettings (lightFieldGrid.cpp:54-58).
void build_subgrid(const float *tri_leaf, uint64_t n, const BvhResult &bvh, uint32_t gs, uint32_t os, uint32_t m, SubgridResult &out);

struct BeamRec { // 48 bytes = 3 x float4, read by subbeam_closest()
    float mn[4], mx[4];
    uint32_t start, end, has_tris, pad;
};
static_assert(sizeof(BeamRec) == 48, "beam record must be 48 bytes");

struct SubbeamResult {
    std::vector<BeamRec> beams;
    std::vector<uint32_t> root[3]; // gridSize^4 root beam ids per dominant axis
    std::vector<uint32_t> idx;     // leaf-order triangle indices
    float gmin[3][3], gmax[3][3], amin[3], amax[3];
    uint32_t grid_size = 0;
};

// scene_lo/hi: scene bounds (getSceneMinBounds/MaxBounds, scene.cpp:42-62).  max_depth guards the
// reference's unbounded recursion; 0 = unbounded.
void build_subbeams(const float *tri_leaf, uint64_t n, const float scene_lo[3], const float scene_hi[3], uint32_t gs, uint32_t m,
                    int max_depth, SubbeamResult &out);

} // namespace ptb
