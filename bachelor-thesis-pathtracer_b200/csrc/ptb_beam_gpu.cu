// ptb_beam_gpu.cu — device executor for the level-synchronous Subbeam builder (ptb_beam_levels.h).
// Stands in for buildGridsWithSubBeams (reference: src/accelerationStructures/lightFieldGridSubBeams.cpp:185-374).
//
// Every step of the driver is one launch over the beams of a depth (one thread per beam, the functor of ptb_beam_levels.h
// as the kernel body), one cub scan, or one cub reduction whose 8-byte result the host reads to size the next level.
// The functors are the ones tests/native/subbeam_levels_check.cpp runs with plain loops against build_subbeams, and this
// file is compiled with the arithmetic contract of the rest of csrc/ (-fmad=false ...), so the pools are the host
// builder's byte for byte (tests/test_gpu_subbeam_build.py compares digests).
#include "ptb_beam_gpu.h"

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <vector>

#include <cub/cub.cuh>

#include "ptb_beam_levels.h"

namespace ptb {
namespace {

constexpr int kBlock = 128;

template <class F>
__global__ void __launch_bounds__(kBlock) k_beam_run(uint32_t n, F f) {
    const uint32_t i = blockIdx.x * (uint32_t)kBlock + threadIdx.x;
    if (i < n) f(i);
}

struct Widen {
    __host__ __device__ unsigned long long operator()(uint32_t v) const { return v; }
};

struct DeviceExec {
    cudaStream_t stream;
    cudaError_t status = cudaSuccess;
    std::vector<void *> scratch;
    char *tmp = nullptr;
    size_t tmpBytes = 0;
    unsigned long long *dSum = nullptr;

    explicit DeviceExec(cudaStream_t s) : stream(s) {}
    ~DeviceExec() {
        cudaStreamSynchronize(stream);
        for (void *p : scratch) cudaFree(p);
        cudaFree(tmp);
        cudaFree(dSum);
    }
    bool ok(cudaError_t e) {
        if (e != cudaSuccess && status == cudaSuccess) status = e;
        return status == cudaSuccess;
    }
    bool fail() const { return status != cudaSuccess; }

    template <class T> T *raw(size_t n) {
        if (fail()) return nullptr;
        void *p = nullptr;
        const size_t bytes = std::max<size_t>(1, n) * sizeof(T);
        if (!ok(cudaMalloc(&p, bytes))) return nullptr;
        if (!ok(cudaMemsetAsync(p, 0, bytes, stream))) {
            cudaFree(p);
            return nullptr;
        }
        return (T *)p;
    }
    template <class T> T *alloc(size_t n) {
        T *p = raw<T>(n);
        if (p) scratch.push_back(p);
        return p;
    }
    template <class T> T *alloc_out(size_t n) { return raw<T>(n); }

    template <class F> void run(uint32_t n, const F &f) {
        if (fail() || n == 0) return; // after an error every pointer may be null: launch nothing, the driver asks fail()
        k_beam_run<F><<<(n + kBlock - 1) / kBlock, kBlock, 0, stream>>>(n, f);
        ok(cudaGetLastError());
    }
    bool need_tmp(size_t bytes) {
        if (bytes <= tmpBytes) return true;
        ok(cudaStreamSynchronize(stream)); // a scan that still reads the old buffer
        cudaFree(tmp);
        tmp = nullptr;
        tmpBytes = 0;
        if (!ok(cudaMalloc(&tmp, bytes))) return false;
        tmpBytes = bytes;
        return true;
    }
    void scan(const uint32_t *in, uint32_t *out, uint32_t n) {
        if (fail() || n == 0) return;
        size_t bytes = 0;
        if (!ok(cub::DeviceScan::ExclusiveSum(nullptr, bytes, in, out, (int)n, stream)) || !need_tmp(bytes)) return;
        ok(cub::DeviceScan::ExclusiveSum(tmp, bytes, in, out, (int)n, stream));
    }
    uint64_t sum64(const uint32_t *in, uint32_t n) {
        if (fail() || n == 0) return 0;
        if (!dSum && !ok(cudaMalloc(&dSum, sizeof *dSum))) return 0;
        cub::TransformInputIterator<unsigned long long, Widen, const uint32_t *> it(in, Widen());
        size_t bytes = 0;
        if (!ok(cub::DeviceReduce::Sum(nullptr, bytes, it, dSum, (int)n, stream)) || !need_tmp(bytes)) return 0;
        if (!ok(cub::DeviceReduce::Sum(tmp, bytes, it, dSum, (int)n, stream))) return 0;
        unsigned long long v = 0;
        if (!ok(cudaMemcpyAsync(&v, dSum, sizeof v, cudaMemcpyDeviceToHost, stream)) || !ok(cudaStreamSynchronize(stream))) return 0;
        return v;
    }
};

} // namespace

cudaError_t subbeams_build_gpu(const float *d_vtx, const uint32_t *d_id_of_tri, uint64_t n, const float scene_lo[3],
                               const float scene_hi[3], uint32_t gs, uint32_t m, int max_depth, cudaStream_t stream, SubbeamDev &out,
                               std::string &err) {
    out = SubbeamDev();
    const uint64_t cells = (uint64_t)gs * gs * gs * gs;
    if (gs < 1 || 3 * cells >= (1ull << 31) || n >= (1ull << 31)) {
        err = "Subbeam grid exceeds 2^31 root beams or triangles";
        return cudaErrorInvalidValue;
    }
    // the enlarged planes per dominant axis, as build_subbeams computes them (ptb_grid_build.cpp; adjustGridSize)
    V3 slo{scene_lo[0], scene_lo[1], scene_lo[2]}, shi{scene_hi[0], scene_hi[1], scene_hi[2]};
    if (n == 0) {
        slo = V3{0, 0, 0};
        shi = V3{0, 0, 0};
    }
    BeamGeom G;
    for (int axis = 0; axis < 3; ++axis) {
        out.amin[axis] = slo[axis];
        out.amax[axis] = shi[axis];
        V3 gmin = slo, gmax = shi;
        enlarge(gmin, gmax, axis);
        for (int k = 0; k < 3; ++k) {
            G.gmin[axis][k] = out.gmin[axis][k] = gmin[k];
            G.gmax[axis][k] = out.gmax[axis][k] = gmax[k];
        }
    }
    G.gs = gs;
    G.cells = (uint32_t)cells;
    G.gsf = (float)gs;
    G.inv = 1.0f / G.gsf;

    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    cudaError_t e = cudaEventCreate(&ev0);
    if (e == cudaSuccess) e = cudaEventCreate(&ev1);
    if (e == cudaSuccess) e = cudaEventRecord(ev0, stream);
    BeamLevelsResult R;
    const char *what = "";
    if (e == cudaSuccess) {
        DeviceExec x(stream);
        // Below the roots, too, one thread per (beam, k entries of the parent's list) instead of one per beam: the same pools
        // (tests/native/subbeam_levels_check.cpp runs both), and on the Lucy stand-in gs 6 / m 50 the build takes 1.40 s instead of
        // 4.09 s at depth guard 2 and 0.50 s instead of 2.06 s at depth 1 with k = 64 / 256 (profiles/r2_subbeam_deep_chunk.jsonl).
        // PTB_BEAM_DEEP_CHUNK=k overrides, 0 = one thread per beam.
        const char *deep = getenv("PTB_BEAM_DEEP_CHUNK");
        const uint32_t deepChunk = deep ? (uint32_t)std::max(0, atoi(deep)) : 64u;
        what = build_beam_levels(x, G, (uint32_t)n, d_vtx, d_id_of_tri, m, max_depth, R, deepChunk);
        e = x.status;
        if (e == cudaSuccess && *what) e = cudaErrorInvalidValue; // a size limit of the driver, nothing wrong with the device
        if (e == cudaSuccess) e = cudaEventRecord(ev1, stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(stream);
    } // scratch freed here
    float ms = 0.0f;
    if (e == cudaSuccess) e = cudaEventElapsedTime(&ms, ev0, ev1);
    if (ev0) cudaEventDestroy(ev0);
    if (ev1) cudaEventDestroy(ev1);
    if (e != cudaSuccess) {
        cudaFree(R.beams);
        cudaFree(R.idx);
        err = *what ? what : std::string("Subbeam build on the device: ") + cudaGetErrorString(e);
        return e;
    }
    out.beams = reinterpret_cast<float4 *>(R.beams);
    out.idx = R.idx;
    out.n_beams = R.nBeams;
    out.n_idx = R.nIdx;
    out.levels = R.depth;
    out.ms = ms;
    return cudaSuccess;
}

} // namespace ptb
