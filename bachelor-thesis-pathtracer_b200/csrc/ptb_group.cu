// ptb_group.cu — the GPUs of one box behind ONE host thread of the reference (include/ptb.h, ptb_group_*).
//
// The reference renders on 16 std::threads that own interleaved pixels (tracer.cpp:228-231); the multi-GPU
// generalisation is SURVEY.md §8e: scene and structures replicated in every GPU's HBM, the frame cut into small
// tiles dealt round-robin, no data-path exchange during the render, one gather of the radiance at the end.
// bench.py does this with one process per GPU and an NCCL reduce (the driver's launch contract); a C++ host —
// the facade that replaces src/tracer.cpp — is one process, so here a group owns one context per device:
//   * ptb_group_render enqueues the path kernel on every device (one launch each, nothing read back) and only
//     then waits, so the GPUs run concurrently under a single host thread;
//   * ptb_group_read_image gathers on device 0 with ONE kernel that reads every pixel straight from its owner's
//     HBM over NVLink (peer loads; each pixel crosses the switch once, 16 B/pixel in total) — owners are known
//     from the tile rule, so no reduction is needed and the frame is bit-identical to the 1-GPU frame.  Where two
//     devices cannot map each other the owner's buffer is staged with cudaMemcpyPeerAsync first.
// Built on the public C ABI only (ptb_image_device_ptr, ptb_stream, ...): a group adds no second rendering path.
#include <cuda_runtime.h>

#include <algorithm>
#include <string>
#include <vector>

#include "ptb.h"

struct ptb_group {
    std::vector<int> dev;
    std::vector<ptb_ctx *> ctx;
    std::vector<char> direct;      // member's image is readable from device 0 (same device or peer access enabled)
    std::vector<float4 *> stage;   // otherwise: a copy on device 0
    size_t stagePixels = 0;
    float4 *dGather = nullptr;     // device 0: the assembled frame
    const float4 **dPtrs = nullptr; // device 0: where to read each member's pixels from
    float *dWeights = nullptr;     // device 0: n_i + 1 per member (PTB_SHARD_SAMPLES)
    size_t gatherPixels = 0;
    ptb_params prm{};
    bool haveParams = false;
    int shard = PTB_SHARD_TILES;
    std::vector<int> samples;      // PTB_SHARD_SAMPLES: samples each member has accumulated since the last reset / clear
    double lastMs = 0.0;
    std::string err;
};

namespace {

std::string g_groupCreateErr;

int gfail(ptb_group *g, int code, const std::string &msg) {
    if (g) g->err = msg;
    else g_groupCreateErr = msg;
    return code;
}

// pixel (x,y) belongs to member ((y / tile_h) * tiles_x + x / tile_w) % n — the rule of alloc_frame (ptb_api.cu)
__global__ void __launch_bounds__(256) k_group_gather(const float4 *const *__restrict__ src, int n, int W, int H, int tileW, int tileH,
                                                       float4 *__restrict__ out) {
    const int tilesX = (W + tileW - 1) / tileW;
    const size_t total = (size_t)W * H;
    for (size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x; p < total; p += (size_t)gridDim.x * blockDim.x) {
        const int y = (int)(p / (size_t)W), x = (int)(p % (size_t)W);
        const int owner = ((y / tileH) * tilesX + x / tileW) % n;
        out[p] = src[owner][p]; // a 16-byte load from the owner's HBM (NVLink when it is a peer)
    }
}

// PTB_SHARD_SAMPLES: member i holds the reference's running mean over its n_i samples, sum(L)/(n_i+1)
// (tracer.cpp:147-154).  Rebuild the sums, add them, apply the same law over all samples: sum(L)/(n_total+1).
__global__ void __launch_bounds__(256) k_group_combine(const float4 *const *__restrict__ src, const float *__restrict__ weight, int n,
                                                        float denom /* n_total + 1 */, size_t total, float4 *__restrict__ out) {
    for (size_t p = (size_t)blockIdx.x * blockDim.x + threadIdx.x; p < total; p += (size_t)gridDim.x * blockDim.x) {
        float4 acc = make_float4(0.0f, 0.0f, 0.0f, 1.0f);
        for (int i = 0; i < n; ++i) {
            const float4 v = src[i][p];
            const float w = weight[i];
            acc.x += v.x * w;
            acc.y += v.y * w;
            acc.z += v.z * w;
        }
        out[p] = make_float4(acc.x / denom, acc.y / denom, acc.z / denom, 1.0f);
    }
}

#define GCK(call)                                                                          \
    do {                                                                                   \
        const cudaError_t e_ = (call);                                                     \
        if (e_ != cudaSuccess) return gfail(g, PTB_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_)); \
    } while (0)

#define GEACH(expr, what)                                                                                  \
    do {                                                                                                   \
        for (size_t i_ = 0; i_ < g->ctx.size(); ++i_) {                                                    \
            ptb_ctx *c = g->ctx[i_];                                                                       \
            const int r_ = (expr);                                                                         \
            if (r_) return gfail(g, r_, std::string(what) + " on member " + std::to_string(i_) + ": " + ptb_last_error(c)); \
        }                                                                                                  \
    } while (0)

void free_gather(ptb_group *g) {
    if (g->dev.empty()) return;
    cudaSetDevice(g->dev[0]);
    cudaFree(g->dGather);
    cudaFree((void *)g->dPtrs);
    cudaFree(g->dWeights);
    g->dWeights = nullptr;
    for (auto *s : g->stage) cudaFree(s);
    g->dGather = nullptr;
    g->dPtrs = nullptr;
    g->stage.assign(g->dev.size(), nullptr);
    g->gatherPixels = g->stagePixels = 0;
}

} // namespace

extern "C" int ptb_group_create(const int *devices, int n_devices, ptb_group **out) {
    if (!out) return PTB_ERR_INVALID;
    *out = nullptr;
    if (!devices || n_devices < 1 || n_devices > 64) return gfail(nullptr, PTB_ERR_INVALID, "ptb_group_create: 1..64 devices");
    ptb_group *g = new ptb_group();
    g->dev.assign(devices, devices + n_devices);
    g->direct.assign(n_devices, 0);
    g->stage.assign(n_devices, nullptr);
    g->samples.assign(n_devices, 0);
    for (int i = 0; i < n_devices; ++i) {
        ptb_ctx *c = nullptr;
        const int r = ptb_create(devices[i], &c);
        if (r) {
            g_groupCreateErr = std::string("ptb_group_create: device ") + std::to_string(devices[i]) + ": " + ptb_last_error(nullptr);
            for (auto *m : g->ctx) ptb_destroy(m);
            delete g;
            return r;
        }
        g->ctx.push_back(c);
    }
    // device 0 reads the other members' frames in place where the hardware lets it (NVLink / NVSwitch peers)
    cudaSetDevice(g->dev[0]);
    for (int i = 0; i < n_devices; ++i) {
        if (g->dev[i] == g->dev[0]) {
            g->direct[i] = 1;
            continue;
        }
        int can = 0;
        if (cudaDeviceCanAccessPeer(&can, g->dev[0], g->dev[i]) == cudaSuccess && can) {
            const cudaError_t e = cudaDeviceEnablePeerAccess(g->dev[i], 0);
            if (e == cudaSuccess || e == cudaErrorPeerAccessAlreadyEnabled) g->direct[i] = 1;
            cudaGetLastError(); // AlreadyEnabled is sticky-free but reported: clear it
        }
    }
    *out = g;
    return PTB_OK;
}

extern "C" void ptb_group_destroy(ptb_group *g) {
    if (!g) return;
    free_gather(g);
    for (auto *c : g->ctx) ptb_destroy(c);
    delete g;
}

extern "C" int ptb_group_size(const ptb_group *g) { return g ? (int)g->ctx.size() : 0; }
extern "C" ptb_ctx *ptb_group_ctx(ptb_group *g, int rank) { return (g && rank >= 0 && rank < (int)g->ctx.size()) ? g->ctx[rank] : nullptr; }
extern "C" const char *ptb_group_last_error(const ptb_group *g) { return g ? g->err.c_str() : g_groupCreateErr.c_str(); }
extern "C" int ptb_group_peer_direct(const ptb_group *g, int rank) { return (g && rank >= 0 && rank < (int)g->ctx.size()) ? g->direct[rank] : 0; }
extern "C" double ptb_group_last_render_ms(const ptb_group *g) { return g ? g->lastMs : 0.0; }

// How the frame is split.  TILES (default): bit-identical to the 1-GPU frame.  SAMPLES (opt-in): every member renders all
// pixels with its share of the samples from its own seeds (ptb_params.seed_offset = rank; member 0 keeps the reference's) —
// the same estimator with other random numbers, statistical parity only; it keeps every GPU full when 1/n of the frame
// has fewer pixels than a GPU has lanes.  Takes effect at the next ptb_group_set_params.
extern "C" int ptb_group_set_sharding(ptb_group *g, int shard) {
    if (!g || (shard != PTB_SHARD_TILES && shard != PTB_SHARD_SAMPLES)) return PTB_ERR_INVALID;
    g->shard = shard;
    g->haveParams = false;
    return PTB_OK;
}

extern "C" int ptb_group_set_params(ptb_group *g, const ptb_params *p) {
    if (!g || !p) return PTB_ERR_INVALID;
    ptb_params q = *p;
    const int n = (int)g->ctx.size();
    if (q.tile_w <= 0) q.tile_w = 32;
    if (q.tile_h <= 0) q.tile_h = 8;
    for (int i = 0; i < n; ++i) {
        if (g->shard == PTB_SHARD_SAMPLES) {
            q.world = 1;
            q.rank = 0;
            q.seed_offset = i;
        } else {
            q.world = n;
            q.rank = i;
        }
        const int r = ptb_set_params(g->ctx[i], &q);
        if (r) return gfail(g, r, std::string("ptb_set_params on member ") + std::to_string(i) + ": " + ptb_last_error(g->ctx[i]));
    }
    q.rank = 0;
    q.world = n;
    g->prm = q;
    g->haveParams = true;
    return PTB_OK;
}

extern "C" int ptb_group_reset(ptb_group *g) {
    if (!g) return PTB_ERR_INVALID;
    GEACH(ptb_reset(c), "ptb_reset");
    std::fill(g->samples.begin(), g->samples.end(), 0);
    return PTB_OK;
}

extern "C" int ptb_group_clear(ptb_group *g) {
    if (!g) return PTB_ERR_INVALID;
    GEACH(ptb_clear(c), "ptb_clear");
    std::fill(g->samples.begin(), g->samples.end(), 0);
    return PTB_OK;
}

// trace(true) on every member: enqueue everywhere first, wait afterwards
extern "C" int ptb_group_render(ptb_group *g, int add_samples) {
    if (!g) return PTB_ERR_INVALID;
    if (!g->haveParams) return gfail(g, PTB_ERR_INVALID, "ptb_group_set_params first");
    if (add_samples <= 0) return gfail(g, PTB_ERR_INVALID, "add_samples must be positive");
    const int n = (int)g->ctx.size();
    for (int i = 0; i < n; ++i) {
        // SAMPLES: member i takes add/n samples, the first add % n members one more
        const int mine = g->shard == PTB_SHARD_SAMPLES ? add_samples / n + (i < add_samples % n ? 1 : 0) : add_samples;
        if (mine <= 0) continue;
        const int r = ptb_render_async(g->ctx[i], mine);
        if (r) return gfail(g, r, std::string("ptb_render_async on member ") + std::to_string(i) + ": " + ptb_last_error(g->ctx[i]));
        g->samples[i] += mine;
    }
    GEACH(ptb_sync(c), "ptb_sync");
    g->lastMs = 0.0;
    for (auto *c : g->ctx) g->lastMs = std::max(g->lastMs, ptb_last_render_ms(c));
    return PTB_OK;
}

extern "C" int ptb_group_read_counters(ptb_group *g, uint64_t out[PTB_CNT__N]) {
    if (!g || !out) return PTB_ERR_INVALID;
    for (int k = 0; k < PTB_CNT__N; ++k) out[k] = 0;
    for (size_t i = 0; i < g->ctx.size(); ++i) {
        uint64_t one[PTB_CNT__N];
        const int r = ptb_read_counters(g->ctx[i], one);
        if (r) return gfail(g, r, std::string("ptb_read_counters on member ") + std::to_string(i) + ": " + ptb_last_error(g->ctx[i]));
        for (int k = 0; k < PTB_CNT__N; ++k) out[k] += one[k];
    }
    return PTB_OK;
}

// tracerGetPixel for every pixel of the whole frame: gathered on device 0 from the owners, then one copy to the host
extern "C" int ptb_group_read_image(ptb_group *g, float *rgba) {
    if (!g || !rgba) return PTB_ERR_INVALID;
    if (!g->haveParams) return gfail(g, PTB_ERR_INVALID, "ptb_group_set_params first");
    const int n = (int)g->ctx.size();
    if (n == 1) {
        const int r = ptb_read_image(g->ctx[0], rgba);
        return r ? gfail(g, r, std::string("ptb_read_image: ") + ptb_last_error(g->ctx[0])) : PTB_OK;
    }
    GEACH(ptb_sync(c), "ptb_sync"); // every member's frame is final
    const size_t px = (size_t)g->prm.width * g->prm.height;
    GCK(cudaSetDevice(g->dev[0]));
    cudaStream_t s0 = (cudaStream_t)ptb_stream(g->ctx[0]);
    if (g->gatherPixels != px) {
        free_gather(g);
        GCK(cudaSetDevice(g->dev[0]));
        GCK(cudaMalloc(&g->dGather, px * sizeof(float4)));
        GCK(cudaMalloc((void **)&g->dPtrs, n * sizeof(float4 *)));
        g->gatherPixels = px;
    }
    std::vector<const float4 *> ptrs(n);
    for (int i = 0; i < n; ++i) {
        const float4 *img = (const float4 *)ptb_image_device_ptr(g->ctx[i]);
        if (!img) return gfail(g, PTB_ERR_INVALID, "member " + std::to_string(i) + " has no frame: ptb_group_reset / ptb_group_render first");
        if (g->direct[i]) {
            ptrs[i] = img;
        } else { // no mapping between the two devices: bring the owner's buffer over first
            if (!g->stage[i] || g->stagePixels != px) {
                cudaFree(g->stage[i]);
                g->stage[i] = nullptr;
                GCK(cudaMalloc(&g->stage[i], px * sizeof(float4)));
            }
            GCK(cudaMemcpyPeerAsync(g->stage[i], g->dev[0], img, g->dev[i], px * sizeof(float4), s0));
            ptrs[i] = g->stage[i];
        }
    }
    g->stagePixels = px;
    GCK(cudaMemcpyAsync((void *)g->dPtrs, ptrs.data(), n * sizeof(float4 *), cudaMemcpyHostToDevice, s0));
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, g->dev[0]);
    const int blocks = (int)std::min<size_t>((px + 255) / 256, (size_t)sms * 8);
    if (g->shard == PTB_SHARD_SAMPLES) {
        std::vector<float> w(n);
        int total = 0;
        for (int i = 0; i < n; ++i) {
            w[i] = g->samples[i] > 0 ? (float)(g->samples[i] + 1) : 0.0f; // a member without samples holds zeros
            total += g->samples[i];
        }
        if (!g->dWeights) GCK(cudaMalloc(&g->dWeights, 64 * sizeof(float)));
        GCK(cudaMemcpyAsync(g->dWeights, w.data(), n * sizeof(float), cudaMemcpyHostToDevice, s0));
        k_group_combine<<<blocks, 256, 0, s0>>>(g->dPtrs, g->dWeights, n, (float)(total + 1), px, g->dGather);
    } else {
        k_group_gather<<<blocks, 256, 0, s0>>>(g->dPtrs, n, g->prm.width, g->prm.height, g->prm.tile_w, g->prm.tile_h, g->dGather);
    }
    GCK(cudaGetLastError());
    GCK(cudaMemcpyAsync(rgba, g->dGather, px * sizeof(float4), cudaMemcpyDeviceToHost, s0));
    GCK(cudaStreamSynchronize(s0));
    return PTB_OK;
}
