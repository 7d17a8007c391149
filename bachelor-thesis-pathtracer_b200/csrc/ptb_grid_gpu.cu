// ptb_grid_gpu.cu — channel population of the Subgrid hierarchy on the GPU (see ptb_grid_gpu.h).
// Stands in for constructChannel / the size^4 loop of constructGrid
// (reference: src/accelerationStructures/lightFieldGrid.cpp:381-414,493-583; src/types/sat.cpp).
//
// One thread per (grid, channel).  The 32 lanes of a warp are consecutive channels of one grid (t fastest), so they
// walk the same candidate list: candidate loads are warp-wide broadcasts and the lanes diverge only where one
// channel finds its separating axis earlier than its neighbour.  Pass 1 keeps a 64-bit accept mask per channel when
// every grid has <= 64 candidates (the published settings: m = 50, 16 children), so pass 2 only expands masks;
// otherwise pass 2 repeats the tests.
#include "ptb_grid_gpu.h"

#include <algorithm>
#include <cub/device/device_scan.cuh>

namespace ptb {
namespace {

constexpr int kBlock = 128;

__device__ __forceinline__ uint32_t grid_of_cell(const uint32_t *__restrict__ cellBase, uint32_t nGrids, uint32_t cell) {
    uint32_t lo = 0, hi = nGrids; // last g with cellBase[g] <= cell
    while (hi - lo > 1) {
        const uint32_t mid = (lo + hi) >> 1;
        if (__ldg(cellBase + mid) <= cell) lo = mid;
        else hi = mid;
    }
    return lo;
}

__global__ void __launch_bounds__(kBlock) k_grid_count(const GridTopo *__restrict__ grids, uint32_t nGrids, const uint32_t *__restrict__ cellBase,
                                                        uint64_t nCells, const uint32_t *__restrict__ cand, const float *__restrict__ vtx,
                                                        const uint32_t *__restrict__ idOfTri, uint32_t *__restrict__ counts,
                                                        unsigned long long *__restrict__ masks, unsigned long long *__restrict__ total) {
    const uint64_t cell = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t cnt = 0;
    if (cell < nCells) {
        const uint32_t g = grid_of_cell(cellBase, nGrids, (uint32_t)cell);
        unsigned long long m = 0ull;
        cnt = populate_cell(grids[g], (uint32_t)cell - __ldg(cellBase + g), grids, cand, vtx, idOfTri, &m, nullptr);
        counts[cell] = cnt;
        if (masks) masks[cell] = m;
    }
    unsigned long long s = cnt;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_down_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0 && s) atomicAdd(total, s);
}

__global__ void __launch_bounds__(kBlock) k_grid_fill(const GridTopo *__restrict__ grids, uint32_t nGrids, const uint32_t *__restrict__ cellBase,
                                                       uint64_t nCells, const uint32_t *__restrict__ cand, const float *__restrict__ vtx,
                                                       const uint32_t *__restrict__ idOfTri, const uint32_t *__restrict__ counts,
                                                       const uint32_t *__restrict__ starts, const unsigned long long *__restrict__ masks,
                                                       uint2 *__restrict__ lut, uint32_t *__restrict__ idx) {
    const uint64_t cell = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (cell >= nCells) return;
    const uint32_t start = __ldg(starts + cell), cnt = __ldg(counts + cell);
    lut[cell] = make_uint2(start, start + cnt);
    if (cnt == 0) return;
    const uint32_t g = grid_of_cell(cellBase, nGrids, (uint32_t)cell);
    if (masks) {
        const GridTopo &G = grids[g];
        const bool hasTris = (G.axis_size_flags >> 31) != 0u;
        unsigned long long m = __ldg(masks + cell);
        uint32_t *o = idx + start;
        while (m) {
            const int k = __ffsll((long long)m) - 1;
            m &= m - 1;
            *o++ = hasTris ? G.cand_base + (uint32_t)k : __ldg(cand + G.cand_base + k);
        }
    } else {
        populate_cell(grids[g], (uint32_t)cell - __ldg(cellBase + g), grids, cand, vtx, idOfTri, nullptr, idx + start);
    }
}

struct Scratch {
    void *p[12] = {nullptr};
    int n = 0;
    template <typename T>
    cudaError_t alloc(T *&out, size_t count) {
        void *q = nullptr;
        const cudaError_t e = cudaMalloc(&q, std::max<size_t>(count, 1) * sizeof(T));
        if (e == cudaSuccess) p[n++] = q;
        out = static_cast<T *>(q);
        return e;
    }
    ~Scratch() {
        for (int i = 0; i < n; ++i) cudaFree(p[i]);
    }
};

void free_dev(SubgridDev &d) {
    cudaFree(d.hdr);
    cudaFree(d.lut);
    cudaFree(d.idx);
    d = SubgridDev();
}

} // namespace

#define GB(call)                                                      \
    do {                                                              \
        const cudaError_t e_ = (call);                                \
        if (e_ != cudaSuccess) {                                      \
            err = std::string(#call) + ": " + cudaGetErrorString(e_); \
            free_dev(out);                                            \
            return e_;                                                \
        }                                                             \
    } while (0)

cudaError_t subgrid_build_gpu(const SubgridTopo &topo, const float *d_vtx, const uint32_t *d_id_of_tri, cudaStream_t stream,
                              SubgridDev &out, std::string &err) {
    out = SubgridDev();
    const size_t nGrids = topo.grids.size();
    const uint64_t nCells = topo.cell_base.back();
    if (nCells >= (1ull << 31) || nGrids >= (1ull << 31)) {
        err = "Subgrid LUT exceeds 2^31 cells";
        return cudaErrorInvalidValue;
    }
    std::vector<GridHdr> hdr(nGrids);
    std::vector<uint32_t> base32(nGrids + 1);
    for (size_t g = 0; g < nGrids; ++g) {
        hdr[g] = subgrid_header(topo.grids[g], topo.cell_base[g]);
        base32[g] = (uint32_t)topo.cell_base[g];
    }
    base32[nGrids] = (uint32_t)nCells;

    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    GB(cudaEventCreate(&ev0));
    GB(cudaEventCreate(&ev1));
    struct EvGuard {
        cudaEvent_t &a, &b;
        ~EvGuard() {
            if (a) cudaEventDestroy(a);
            if (b) cudaEventDestroy(b);
        }
    } evGuard{ev0, ev1};

    Scratch S;
    GridTopo *dGrids = nullptr;
    uint32_t *dBase = nullptr, *dCand = nullptr, *dCounts = nullptr, *dStarts = nullptr;
    unsigned long long *dMasks = nullptr, *dTotal = nullptr;
    char *tmp = nullptr;
    const bool useMasks = topo.max_cand <= 64;
    GB(S.alloc(dGrids, nGrids));
    GB(S.alloc(dBase, nGrids + 1));
    GB(S.alloc(dCand, topo.cand.size()));
    GB(S.alloc(dCounts, nCells));
    GB(S.alloc(dStarts, nCells));
    GB(S.alloc(dTotal, 1));
    if (useMasks) GB(S.alloc(dMasks, nCells));
    size_t scanBytes = 0;
    GB(cub::DeviceScan::ExclusiveSum(nullptr, scanBytes, dCounts, dStarts, (int)nCells, stream));
    GB(S.alloc(tmp, scanBytes));
    GB(cudaMalloc(&out.hdr, std::max<size_t>(1, nGrids) * sizeof(GridHdr)));
    GB(cudaMalloc(&out.lut, std::max<uint64_t>(1, nCells) * sizeof(uint2)));
    GB(cudaMemcpyAsync(dGrids, topo.grids.data(), nGrids * sizeof(GridTopo), cudaMemcpyHostToDevice, stream));
    GB(cudaMemcpyAsync(dBase, base32.data(), base32.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, stream));
    if (!topo.cand.empty()) GB(cudaMemcpyAsync(dCand, topo.cand.data(), topo.cand.size() * sizeof(uint32_t), cudaMemcpyHostToDevice, stream));
    GB(cudaMemcpyAsync(out.hdr, hdr.data(), nGrids * sizeof(GridHdr), cudaMemcpyHostToDevice, stream));
    GB(cudaMemsetAsync(dTotal, 0, sizeof(unsigned long long), stream));

    const unsigned blocks = (unsigned)((nCells + kBlock - 1) / kBlock);
    GB(cudaEventRecord(ev0, stream));
    unsigned long long total = 0;
    if (nCells) {
        k_grid_count<<<blocks, kBlock, 0, stream>>>(dGrids, (uint32_t)nGrids, dBase, nCells, dCand, d_vtx, d_id_of_tri, dCounts, dMasks, dTotal);
        GB(cudaGetLastError());
        GB(cub::DeviceScan::ExclusiveSum(tmp, scanBytes, dCounts, dStarts, (int)nCells, stream));
        GB(cudaMemcpyAsync(&total, dTotal, sizeof total, cudaMemcpyDeviceToHost, stream));
        GB(cudaStreamSynchronize(stream));
        if (total >= (1ull << 32)) {
            err = "Subgrid pools exceed 2^32 entries";
            free_dev(out);
            return cudaErrorInvalidValue;
        }
    }
    GB(cudaMalloc(&out.idx, std::max<uint64_t>(1, total) * sizeof(uint32_t)));
    if (nCells) {
        k_grid_fill<<<blocks, kBlock, 0, stream>>>(dGrids, (uint32_t)nGrids, dBase, nCells, dCand, d_vtx, d_id_of_tri, dCounts, dStarts, dMasks,
                                                   out.lut, out.idx);
        GB(cudaGetLastError());
    }
    GB(cudaEventRecord(ev1, stream));
    GB(cudaStreamSynchronize(stream));
    float ms = 0.0f;
    GB(cudaEventElapsedTime(&ms, ev0, ev1));
    out.ms = ms;
    out.n_grids = nGrids;
    out.n_cells = nCells;
    out.n_idx = total;
    return cudaSuccess;
}

} // namespace ptb
