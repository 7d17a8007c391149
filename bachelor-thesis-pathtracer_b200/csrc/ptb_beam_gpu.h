// ptb_beam_gpu.h — the Subbeam structure built on the device (second half): ptb_beam_levels.h's
// level-synchronous driver with CUDA kernels and cub behind it.
// Stands in for buildGridsWithSubBeams (reference: src/accelerationStructures/lightFieldGridSubBeams.cpp:185-374).
#pragma once
#include <cstdint>
#include <string>

#include <cuda_runtime.h>

namespace ptb {

struct SubbeamDev { // device pools, cudaMalloc'ed by subbeams_build_gpu; the caller owns (cudaFree) them
    float4 *beams = nullptr; // 48 B per beam (BeamRec): roots at [axis * gs^4 + cell], descendants appended per tree
    uint32_t *idx = nullptr; // leaf-order triangle ids of the leaf beams
    uint64_t n_beams = 0, n_idx = 0;
    float gmin[3][3], gmax[3][3], amin[3], amax[3];
    int levels = 0;  // deepest level that holds beams
    double ms = 0.0; // stream time of the build (CUDA events; includes the host's per-level size read-backs)
};

// d_vtx: 9 floats per triangle by reference id, d_id_of_tri: leaf position -> reference id (device pointers).
// scene_lo/hi: getSceneMinBounds/MaxBounds.  max_depth: the guard of ptb_set_build_limits (0 = none).
cudaError_t subbeams_build_gpu(const float *d_vtx, const uint32_t *d_id_of_tri, uint64_t n, const float scene_lo[3],
                               const float scene_hi[3], uint32_t gs, uint32_t m, int max_depth, cudaStream_t stream, SubbeamDev &out,
                               std::string &err);

} // namespace ptb
