// ptb_lbvh.cu — the BVH built in HBM (see ptb_lbvh.h).  Stands in for constructBVH
// (reference: src/accelerationStructures/bvh.cpp:222-390) where the host builders stop scaling.
//
// Pipeline, every stage one kernel over flat arrays (HBM-bound streams and gathers, grids sized in
// multiples of the SM count):
//   k_bounds     scene box: per-thread min/max -> warp shuffle -> 6 ordered-int atomics per warp
//   k_morton     63-bit Morton key of each triangle's box centre (21 bits per axis; fewer on a depth retry)
//   cub radix sort of (key, reference id)
//   k_hierarchy  Karras' binary radix tree: node i finds its key range and split by binary search on the
//                common-prefix length (ties between equal keys broken by the sorted position)
//   k_refit      bottom-up boxes: one thread per triangle climbs; the second arrival at a node (atomic
//                counter) joins its own box with the sibling's and goes on, the first one retires
//   k_kept + cub exclusive scan: subtrees of <= leaf_max triangles collapse into one leaf (their triangles
//                are contiguous in sorted order), the remaining inner nodes get dense indices
//   k_emit       the 64-byte two-box nodes the traversal reads (BvhNode64)
//   k_gather     leaf-ordered (v0, e1, e2) / normals / id maps; edges pre-subtracted in fp32 exactly like
//                triangle.cpp:10-13 (the library is compiled with -fmad=false)
#include "ptb_lbvh.h"

#include <algorithm>
#include <cfloat>
#include <climits>
#include <cstring>
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

namespace ptb {
namespace {

constexpr int kBlock = 256;
constexpr uint32_t kLeafFlag = 0x80000000u; // child reference inside the radix tree: leaf position | flag

// order-preserving float <-> int map for atomicMin/atomicMax
__device__ __forceinline__ int f2ord(float f) {
    const int b = __float_as_int(f);
    return b >= 0 ? b : b ^ 0x7fffffff;
}
__host__ __device__ __forceinline__ float ord2f(int o) {
    const int b = o >= 0 ? o : o ^ 0x7fffffff;
#ifdef __CUDA_ARCH__
    return __int_as_float(b);
#else
    float f;
    memcpy(&f, &b, 4);
    return f;
#endif
}

struct Box3 {
    float lx, ly, lz, hx, hy, hz;
};

__device__ __forceinline__ Box3 tri_box(const float *__restrict__ vtx, uint32_t id) {
    const float *p = vtx + 9 * (size_t)id;
    float v[9];
#pragma unroll
    for (int k = 0; k < 9; ++k) v[k] = __ldg(p + k);
    Box3 b;
    b.lx = fminf(v[0], fminf(v[3], v[6]));
    b.ly = fminf(v[1], fminf(v[4], v[7]));
    b.lz = fminf(v[2], fminf(v[5], v[8]));
    b.hx = fmaxf(v[0], fmaxf(v[3], v[6]));
    b.hy = fmaxf(v[1], fmaxf(v[4], v[7]));
    b.hz = fmaxf(v[2], fmaxf(v[5], v[8]));
    return b;
}

__device__ __forceinline__ void join(Box3 &a, const Box3 &b) {
    a.lx = fminf(a.lx, b.lx);
    a.ly = fminf(a.ly, b.ly);
    a.lz = fminf(a.lz, b.lz);
    a.hx = fmaxf(a.hx, b.hx);
    a.hy = fmaxf(a.hy, b.hy);
    a.hz = fmaxf(a.hz, b.hz);
}

// ---- scene bounds -----------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlock) k_bounds(const float *__restrict__ vtx, uint32_t n, int *__restrict__ ord6) {
    Box3 b{FLT_MAX, FLT_MAX, FLT_MAX, -FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) join(b, tri_box(vtx, i));
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        b.lx = fminf(b.lx, __shfl_xor_sync(0xffffffffu, b.lx, o));
        b.ly = fminf(b.ly, __shfl_xor_sync(0xffffffffu, b.ly, o));
        b.lz = fminf(b.lz, __shfl_xor_sync(0xffffffffu, b.lz, o));
        b.hx = fmaxf(b.hx, __shfl_xor_sync(0xffffffffu, b.hx, o));
        b.hy = fmaxf(b.hy, __shfl_xor_sync(0xffffffffu, b.hy, o));
        b.hz = fmaxf(b.hz, __shfl_xor_sync(0xffffffffu, b.hz, o));
    }
    if ((threadIdx.x & 31) == 0) {
        atomicMin(ord6 + 0, f2ord(b.lx));
        atomicMin(ord6 + 1, f2ord(b.ly));
        atomicMin(ord6 + 2, f2ord(b.lz));
        atomicMax(ord6 + 3, f2ord(b.hx));
        atomicMax(ord6 + 4, f2ord(b.hy));
        atomicMax(ord6 + 5, f2ord(b.hz));
    }
}

// ---- Morton keys ------------------------------------------------------------------------------
__device__ __forceinline__ uint64_t spread21(uint32_t v) { // 21 bits -> every third bit of 63
    uint64_t x = v & 0x1fffffu;
    x = (x | x << 32) & 0x1f00000000ffffull;
    x = (x | x << 16) & 0x1f0000ff0000ffull;
    x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

__device__ __forceinline__ uint32_t quant21(float c, float lo, float inv) {
    float q = (c - lo) * inv * 2097152.0f;
    q = fminf(fmaxf(q, 0.0f), 2097151.0f);
    return (uint32_t)q;
}

__global__ void __launch_bounds__(kBlock) k_morton(const float *__restrict__ vtx, uint32_t n, const int *__restrict__ ord6, int dropBits,
                                                    uint64_t *__restrict__ keys, uint32_t *__restrict__ vals) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float lx = ord2f(ord6[0]), ly = ord2f(ord6[1]), lz = ord2f(ord6[2]);
    const float ex = ord2f(ord6[3]) - lx, ey = ord2f(ord6[4]) - ly, ez = ord2f(ord6[5]) - lz;
    const float ix = ex > 0.0f ? 1.0f / ex : 0.0f, iy = ey > 0.0f ? 1.0f / ey : 0.0f, iz = ez > 0.0f ? 1.0f / ez : 0.0f;
    const Box3 b = tri_box(vtx, i);
    const uint32_t qx = quant21(0.5f * (b.lx + b.hx), lx, ix) >> dropBits;
    const uint32_t qy = quant21(0.5f * (b.ly + b.hy), ly, iy) >> dropBits;
    const uint32_t qz = quant21(0.5f * (b.lz + b.hz), lz, iz) >> dropBits;
    keys[i] = (spread21(qx) << 2) | (spread21(qy) << 1) | spread21(qz);
    vals[i] = i;
}

// ---- radix tree (Karras 2012) -----------------------------------------------------------------
// common-prefix length of the keys at sorted positions i and j; equal keys are told apart by position
__device__ __forceinline__ int delta(const uint64_t *__restrict__ keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    const uint64_t x = __ldg(keys + i) ^ __ldg(keys + j);
    return x ? __clzll((long long)x) : 64 + __clz(i ^ j);
}

__global__ void __launch_bounds__(kBlock) k_hierarchy(const uint64_t *__restrict__ keys, int n, uint32_t *__restrict__ left,
                                                       uint32_t *__restrict__ right, uint32_t *__restrict__ parent,
                                                       uint32_t *__restrict__ leafParent, uint2 *__restrict__ range) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const int d = (delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    const int dmin = delta(keys, n, i, i - d);
    int lmax = 2;
    while (delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = delta(keys, n, i, j);
    int s = 0, t = l;
    do {
        t = (t + 1) >> 1;
        if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    const int gamma = i + s * d + min(d, 0);
    const int lo = min(i, j), hi = max(i, j);
    const uint32_t lref = (lo == gamma) ? ((uint32_t)gamma | kLeafFlag) : (uint32_t)gamma;
    const uint32_t rref = (hi == gamma + 1) ? ((uint32_t)(gamma + 1) | kLeafFlag) : (uint32_t)(gamma + 1);
    left[i] = lref;
    right[i] = rref;
    range[i] = make_uint2((uint32_t)lo, (uint32_t)hi);
    if (lref & kLeafFlag) leafParent[gamma] = (uint32_t)i;
    else parent[gamma] = (uint32_t)i;
    if (rref & kLeafFlag) leafParent[gamma + 1] = (uint32_t)i;
    else parent[gamma + 1] = (uint32_t)i;
}

// ---- bottom-up boxes ---------------------------------------------------------------------------
// boxLo[i] = (lo.xyz, height as int bits), boxHi[i] = (hi.xyz, 0).  height = inner nodes that survive the
// collapse on the longest chain below and including i (what the traversal stack has to hold).
__global__ void __launch_bounds__(kBlock) k_refit(const float *__restrict__ vtx, const uint32_t *__restrict__ vals, uint32_t n,
                                                   const uint32_t *__restrict__ left, const uint32_t *__restrict__ right,
                                                   const uint32_t *__restrict__ parent, const uint32_t *__restrict__ leafParent,
                                                   const uint2 *__restrict__ range, uint32_t leafMax, uint32_t *__restrict__ arrivals,
                                                   float4 *boxLo, float4 *boxHi) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    Box3 b = tri_box(vtx, __ldg(vals + k));
    int h = 0;
    uint32_t me = k | kLeafFlag;
    uint32_t cur = __ldg(leafParent + k);
    for (;;) {
        __threadfence(); // my subtree's box is visible before I announce it
        if (atomicAdd(arrivals + cur, 1u) == 0u) return; // the sibling subtree is not finished: its last thread goes on
        const uint32_t l = __ldg(left + cur), r = __ldg(right + cur);
        const uint32_t sib = (l == me) ? r : l;
        Box3 sb;
        int sh = 0;
        if (sib & kLeafFlag) {
            sb = tri_box(vtx, __ldg(vals + (sib & ~kLeafFlag)));
        } else {
            const float4 a = __ldcg(boxLo + sib), c = __ldcg(boxHi + sib);
            sb = Box3{a.x, a.y, a.z, c.x, c.y, c.z};
            sh = __float_as_int(a.w);
        }
        join(b, sb);
        const uint2 rg = __ldg(range + cur);
        h = (rg.y - rg.x + 1u > leafMax) ? 1 + max(h, sh) : 0;
        boxLo[cur] = make_float4(b.lx, b.ly, b.lz, __int_as_float(h));
        boxHi[cur] = make_float4(b.hx, b.hy, b.hz, 0.0f);
        if (cur == 0u) return;
        me = cur;
        cur = __ldg(parent + cur);
    }
}

// ---- collapse + emit ---------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlock) k_kept(const uint2 *__restrict__ range, uint32_t nInner, uint32_t leafMax,
                                                  uint32_t *__restrict__ kept) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nInner) return;
    const uint2 rg = __ldg(range + i);
    kept[i] = (rg.y - rg.x + 1u > leafMax) ? 1u : 0u;
}

__device__ __forceinline__ int leaf_ref(uint32_t first, uint32_t count) { return (int)~((first << 4) | (count - 1u)); }

__global__ void __launch_bounds__(kBlock) k_emit(const float *__restrict__ vtx, const uint32_t *__restrict__ vals, uint32_t nInner,
                                                  const uint32_t *__restrict__ left, const uint32_t *__restrict__ right,
                                                  const uint2 *__restrict__ range, uint32_t leafMax, const uint32_t *__restrict__ dense,
                                                  const float4 *__restrict__ boxLo, const float4 *__restrict__ boxHi,
                                                  float4 *__restrict__ nodes) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nInner) return;
    const uint2 rg = __ldg(range + i);
    if (rg.y - rg.x + 1u <= leafMax) return; // inside a collapsed leaf
    Box3 bx[2];
    int ref[2];
#pragma unroll
    for (int s = 0; s < 2; ++s) {
        const uint32_t c = s ? __ldg(right + i) : __ldg(left + i);
        if (c & kLeafFlag) {
            const uint32_t k = c & ~kLeafFlag;
            bx[s] = tri_box(vtx, __ldg(vals + k));
            ref[s] = leaf_ref(k, 1u);
        } else {
            const float4 a = __ldg(boxLo + c), b = __ldg(boxHi + c);
            bx[s] = Box3{a.x, a.y, a.z, b.x, b.y, b.z};
            const uint2 cr = __ldg(range + c);
            const uint32_t cnt = cr.y - cr.x + 1u;
            ref[s] = cnt > leafMax ? (int)__ldg(dense + c) : leaf_ref(cr.x, cnt);
        }
    }
    float4 *o = nodes + 4 * (size_t)__ldg(dense + i);
    o[0] = make_float4(bx[0].lx, bx[0].ly, bx[0].lz, bx[0].hx);
    o[1] = make_float4(bx[0].hy, bx[0].hz, bx[1].lx, bx[1].ly);
    o[2] = make_float4(bx[1].lz, bx[1].hx, bx[1].hy, bx[1].hz);
    o[3] = make_float4(__int_as_float(ref[0]), __int_as_float(ref[1]), 0.0f, 0.0f);
}

__global__ void __launch_bounds__(kBlock) k_gather(const float *__restrict__ vtx, const float *__restrict__ nrm9,
                                                    const int32_t *__restrict__ mat, const uint32_t *__restrict__ vals, uint32_t n,
                                                    float4 *__restrict__ tri, float4 *__restrict__ nrm, uint32_t *__restrict__ triOfId,
                                                    uint32_t *__restrict__ idOfTri) {
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n) return;
    const uint32_t id = __ldg(vals + k);
    const float *p = vtx + 9 * (size_t)id, *q = nrm9 + 9 * (size_t)id;
    float v[9], m[9];
#pragma unroll
    for (int j = 0; j < 9; ++j) {
        v[j] = __ldg(p + j);
        m[j] = __ldg(q + j);
    }
    float4 *t = tri + 3 * (size_t)k, *o = nrm + 3 * (size_t)k;
    t[0] = make_float4(v[0], v[1], v[2], __int_as_float((int)id));
    t[1] = make_float4(v[3] - v[0], v[4] - v[1], v[5] - v[2], 0.0f); // edge1 = v1 - v0 (triangle.cpp:10-11)
    t[2] = make_float4(v[6] - v[0], v[7] - v[1], v[8] - v[2], 0.0f); // edge2 = v2 - v0 (triangle.cpp:12-13)
    o[0] = make_float4(m[0], m[1], m[2], __int_as_float(__ldg(mat + id)));
    o[1] = make_float4(m[3], m[4], m[5], 0.0f);
    o[2] = make_float4(m[6], m[7], m[8], 0.0f);
    triOfId[id] = k;
    idOfTri[k] = id;
}

struct Scratch { // frees whatever is still held when it goes out of scope
    void *p[16] = {nullptr};
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

inline unsigned blocks_for(uint64_t n) { return (unsigned)((n + kBlock - 1) / kBlock); }

} // namespace

#define LB(call)                                                           \
    do {                                                                   \
        const cudaError_t e_ = (call);                                     \
        if (e_ != cudaSuccess) {                                           \
            err = std::string(#call) + ": " + cudaGetErrorString(e_);      \
            free_result(out);                                              \
            return e_;                                                     \
        }                                                                  \
    } while (0)

static void free_result(LbvhResult &r) {
    cudaFree(r.tri);
    cudaFree(r.nrm);
    cudaFree(r.nodes);
    cudaFree(r.tri_of_id);
    cudaFree(r.id_of_tri);
    r = LbvhResult();
}

static cudaError_t lbvh_attempt(const float *d_vtx, const float *d_nrm9, const int32_t *d_mat, uint64_t n64, int leaf_max,
                                int bitsPerAxis, cudaStream_t stream, LbvhResult &out, std::string &err) {
    out = LbvhResult();
    if (leaf_max < 1 || leaf_max > 16 || n64 >= (1ull << 27)) {
        err = "lbvh_build: leaf_max must be 1..16 and n < 2^27";
        return cudaErrorInvalidValue;
    }
    const uint32_t n = (uint32_t)n64;
    LB(cudaMalloc(&out.nodes, 64)); // never null: kernels take the pointer even for leaf-only scenes
    if (n == 0) return cudaSuccess;
    const uint32_t nInner = n - 1;
    const uint32_t leafMax = (uint32_t)leaf_max;

    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    LB(cudaEventCreate(&ev0));
    LB(cudaEventCreate(&ev1));
    struct EvGuard {
        cudaEvent_t &a, &b;
        ~EvGuard() {
            if (a) cudaEventDestroy(a);
            if (b) cudaEventDestroy(b);
        }
    } evGuard{ev0, ev1};

    Scratch S;
    int *ord6 = nullptr;
    uint64_t *keys[2] = {nullptr, nullptr};
    uint32_t *vals[2] = {nullptr, nullptr};
    LB(S.alloc(ord6, 6));
    LB(S.alloc(keys[0], n));
    LB(S.alloc(keys[1], n));
    LB(S.alloc(vals[0], n));
    LB(S.alloc(vals[1], n));
    LB(cudaMalloc(&out.tri, 3 * (size_t)n * sizeof(float4)));
    LB(cudaMalloc(&out.nrm, 3 * (size_t)n * sizeof(float4)));
    LB(cudaMalloc(&out.tri_of_id, (size_t)n * sizeof(uint32_t)));
    LB(cudaMalloc(&out.id_of_tri, (size_t)n * sizeof(uint32_t)));

    LB(cudaEventRecord(ev0, stream));
    const int ordInit[6] = {INT_MAX, INT_MAX, INT_MAX, INT_MIN, INT_MIN, INT_MIN};
    LB(cudaMemcpyAsync(ord6, ordInit, sizeof ordInit, cudaMemcpyHostToDevice, stream));
    int sms = 148;
    {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    }
    k_bounds<<<std::min<unsigned>(blocks_for(n), (unsigned)sms * 8u), kBlock, 0, stream>>>(d_vtx, n, ord6);
    k_morton<<<blocks_for(n), kBlock, 0, stream>>>(d_vtx, n, ord6, 21 - bitsPerAxis, keys[0], vals[0]);
    LB(cudaGetLastError());

    cub::DoubleBuffer<uint64_t> dk(keys[0], keys[1]);
    cub::DoubleBuffer<uint32_t> dv(vals[0], vals[1]);
    size_t tmpBytes = 0, scanBytes = 0;
    LB(cub::DeviceRadixSort::SortPairs(nullptr, tmpBytes, dk, dv, (int)n, 0, 3 * bitsPerAxis, stream));
    uint32_t *kept = nullptr, *dense = nullptr;
    LB(S.alloc(kept, nInner));
    LB(S.alloc(dense, nInner));
    LB(cub::DeviceScan::ExclusiveSum(nullptr, scanBytes, kept, dense, (int)std::max(nInner, 1u), stream));
    char *tmp = nullptr;
    LB(S.alloc(tmp, std::max(tmpBytes, scanBytes)));
    LB(cub::DeviceRadixSort::SortPairs(tmp, tmpBytes, dk, dv, (int)n, 0, 3 * bitsPerAxis, stream));
    const uint64_t *sortedKeys = dk.Current();
    const uint32_t *sortedVals = dv.Current();

    int ordHost[6];
    LB(cudaMemcpyAsync(ordHost, ord6, sizeof ordHost, cudaMemcpyDeviceToHost, stream));

    if (nInner > 0) {
        uint32_t *left = nullptr, *right = nullptr, *parent = nullptr, *leafParent = nullptr, *arrivals = nullptr;
        uint2 *range = nullptr;
        float4 *boxLo = nullptr, *boxHi = nullptr;
        LB(S.alloc(left, nInner));
        LB(S.alloc(right, nInner));
        LB(S.alloc(parent, nInner));
        LB(S.alloc(leafParent, n));
        LB(S.alloc(arrivals, nInner));
        LB(S.alloc(range, nInner));
        LB(S.alloc(boxLo, nInner));
        LB(S.alloc(boxHi, nInner));
        LB(cudaMemsetAsync(arrivals, 0, (size_t)nInner * sizeof(uint32_t), stream));
        k_hierarchy<<<blocks_for(nInner), kBlock, 0, stream>>>(sortedKeys, (int)n, left, right, parent, leafParent, range);
        k_refit<<<blocks_for(n), kBlock, 0, stream>>>(d_vtx, sortedVals, n, left, right, parent, leafParent, range, leafMax, arrivals,
                                                      boxLo, boxHi);
        k_kept<<<blocks_for(nInner), kBlock, 0, stream>>>(range, nInner, leafMax, kept);
        LB(cudaGetLastError());
        LB(cub::DeviceScan::ExclusiveSum(tmp, scanBytes, kept, dense, (int)nInner, stream));
        // number of surviving inner nodes = dense[last] + kept[last]; the root's height rides in boxLo[0].w
        uint32_t tail[2] = {0, 0};
        float4 rootLo;
        LB(cudaMemcpyAsync(&tail[0], dense + (nInner - 1), 4, cudaMemcpyDeviceToHost, stream));
        LB(cudaMemcpyAsync(&tail[1], kept + (nInner - 1), 4, cudaMemcpyDeviceToHost, stream));
        LB(cudaMemcpyAsync(&rootLo, boxLo, sizeof rootLo, cudaMemcpyDeviceToHost, stream));
        LB(cudaStreamSynchronize(stream));
        out.n_nodes = (uint64_t)tail[0] + tail[1];
        memcpy(&out.height, &rootLo.w, 4);
        if (out.n_nodes > 0) {
            cudaFree(out.nodes);
            out.nodes = nullptr;
            LB(cudaMalloc(&out.nodes, out.n_nodes * 64));
            k_emit<<<blocks_for(nInner), kBlock, 0, stream>>>(d_vtx, sortedVals, nInner, left, right, range, leafMax, dense, boxLo, boxHi,
                                                               out.nodes);
            out.root = 0; // the radix tree's root is node 0 and the scan is exclusive
        } else {
            out.root = (int32_t)~(((uint32_t)0 << 4) | (n - 1u)); // the whole scene is one leaf
        }
    } else {
        out.root = (int32_t)~0u; // a single triangle: leaf {first 0, count 1}
    }
    k_gather<<<blocks_for(n), kBlock, 0, stream>>>(d_vtx, d_nrm9, d_mat, sortedVals, n, out.tri, out.nrm, out.tri_of_id, out.id_of_tri);
    LB(cudaGetLastError());
    LB(cudaEventRecord(ev1, stream));
    LB(cudaStreamSynchronize(stream));
    float ms = 0.0f;
    LB(cudaEventElapsedTime(&ms, ev0, ev1));
    out.ms = ms;
    for (int k = 0; k < 3; ++k) {
        out.lo[k] = ord2f(ordHost[k]);
        out.hi[k] = ord2f(ordHost[3 + k]);
    }
    return cudaSuccess;
}

// A radix tree over b-bit keys is at most b + log2(longest run of equal keys) deep.  Real scenes stay far below the
// 63 levels full keys allow (Lucy stand-in: see tests), but a geometric cascade of clusters can use them all; then
// coarser keys bound the depth (equal keys split by position, i.e. balanced) at the price of a looser tree.
cudaError_t lbvh_build(const float *d_vtx, const float *d_nrm9, const int32_t *d_mat, uint64_t n, int leaf_max, int max_height,
                       cudaStream_t stream, LbvhResult &out, std::string &err) {
    cudaError_t e = cudaSuccess;
    for (int bits : {21, 14, 10, 6}) {
        e = lbvh_attempt(d_vtx, d_nrm9, d_mat, n, leaf_max, bits, stream, out, err);
        if (e != cudaSuccess || out.height < max_height) return e;
        if (bits != 6) free_result(out);
    }
    return e; // still too deep: the caller reports it
}

} // namespace ptb
