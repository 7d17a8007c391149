// ptb_bvh_build.cpp — binned-SAH BVH2 builder (host), see ptb_bvh_build.h.
#include "ptb_bvh_build.h"

#include <algorithm>
#include <atomic>
#include <cfloat>
#include <cstdlib>
#include <cmath>
#include <cstring>
#include <future>
#include <thread>

namespace ptb {
namespace {

struct Box {
    float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    void grow(const float *p) {
        for (int k = 0; k < 3; ++k) {
            lo[k] = std::min(lo[k], p[k]);
            hi[k] = std::max(hi[k], p[k]);
        }
    }
    void grow(const Box &b) {
        for (int k = 0; k < 3; ++k) {
            lo[k] = std::min(lo[k], b.lo[k]);
            hi[k] = std::max(hi[k], b.hi[k]);
        }
    }
    float half_area() const {
        float ex = hi[0] - lo[0], ey = hi[1] - lo[1], ez = hi[2] - lo[2];
        if (ex < 0 || ey < 0 || ez < 0) return 0.0f;
        return ex * (ey + ez) + ey * ez;
    }
};

struct Builder {
    const float *pos9;
    uint64_t n;
    int leaf_max;
    int bins = 16;
    std::vector<Box> tbox;     // per reference triangle
    std::vector<float> cent;   // 3 per reference triangle
    std::vector<uint32_t> idx; // working order
    std::vector<BvhNode64> nodes;
    std::vector<BvhRange> ranges;
    std::atomic<uint32_t> next{0};
    std::atomic<int> max_depth{0};
    std::atomic<int> live_threads{1};
    int hw_threads = 1;

    static int32_t leaf_ref(uint32_t first, uint32_t count) { return (int32_t)~((first << 4) | (count - 1u)); }

    // builds the subtree over idx[first, first+count); returns its reference and its box
    int32_t build(uint32_t first, uint32_t count, int depth, Box &box) {
        box = Box();
        Box cb;
        for (uint32_t i = first; i < first + count; ++i) {
            box.grow(tbox[idx[i]]);
            cb.grow(&cent[3 * (size_t)idx[i]]);
        }
        int d = max_depth.load(std::memory_order_relaxed);
        while (depth > d && !max_depth.compare_exchange_weak(d, depth)) {
        }
        if ((int)count <= leaf_max) return leaf_ref(first, count);

        // binned SAH sweep on the three axes (centroid binning); `bins` = 16 unless PTB_BVH_BINS says otherwise
        const int NBMAX = 256;
        const int NB = bins;
        int bestAxis = -1, bestBin = -1;
        float bestCost = FLT_MAX;
        if (depth < 32) {
            for (int a = 0; a < 3; ++a) {
                const float lo = cb.lo[a], ext = cb.hi[a] - cb.lo[a];
                if (!(ext > 0.0f)) continue;
                Box bb[NBMAX];
                uint32_t bc[NBMAX] = {0};
                const float scale = (float)NB / ext;
                for (uint32_t i = first; i < first + count; ++i) {
                    const uint32_t t = idx[i];
                    int b = (int)((cent[3 * (size_t)t + a] - lo) * scale);
                    b = b < 0 ? 0 : (b >= NB ? NB - 1 : b);
                    bb[b].grow(tbox[t]);
                    bc[b]++;
                }
                float rightArea[NBMAX];
                uint32_t rightCnt[NBMAX];
                Box acc;
                uint32_t c = 0;
                for (int b = NB - 1; b > 0; --b) {
                    acc.grow(bb[b]);
                    c += bc[b];
                    rightArea[b] = acc.half_area();
                    rightCnt[b] = c;
                }
                acc = Box();
                c = 0;
                for (int b = 0; b < NB - 1; ++b) {
                    acc.grow(bb[b]);
                    c += bc[b];
                    if (c == 0 || rightCnt[b + 1] == 0) continue;
                    const float cost = acc.half_area() * (float)c + rightArea[b + 1] * (float)rightCnt[b + 1];
                    if (cost < bestCost) {
                        bestCost = cost;
                        bestAxis = a;
                        bestBin = b;
                    }
                }
            }
        }
        uint32_t mid;
        if (bestAxis >= 0) {
            const float lo = cb.lo[bestAxis], scale = (float)NB / (cb.hi[bestAxis] - cb.lo[bestAxis]);
            const int a = bestAxis, bin = bestBin;
            auto it = std::partition(idx.begin() + first, idx.begin() + first + count, [&](uint32_t t) {
                int b = (int)((cent[3 * (size_t)t + a] - lo) * scale);
                b = b < 0 ? 0 : (b >= NB ? NB - 1 : b);
                return b <= bin;
            });
            mid = (uint32_t)(it - idx.begin());
        } else {
            // no usable SAH split (coincident centroids or depth guard): object median on the widest axis
            int a = 0;
            float e = -1.0f;
            for (int k = 0; k < 3; ++k)
                if (cb.hi[k] - cb.lo[k] > e) {
                    e = cb.hi[k] - cb.lo[k];
                    a = k;
                }
            mid = first + count / 2;
            std::nth_element(idx.begin() + first, idx.begin() + mid, idx.begin() + first + count,
                             [&](uint32_t x, uint32_t y) { return cent[3 * (size_t)x + a] < cent[3 * (size_t)y + a]; });
        }
        if (mid == first || mid == first + count) mid = first + count / 2;

        const uint32_t me = next.fetch_add(1);
        Box b0, b1;
        int32_t c0, c1;
        const uint32_t nl = mid - first, nr = first + count - mid;
        bool spawned = false;
        if (count > 200000 && live_threads.load() < hw_threads) {
            live_threads.fetch_add(1);
            spawned = true;
            auto fut = std::async(std::launch::async, [&]() { return build(first, nl, depth + 1, b0); });
            c1 = build(mid, nr, depth + 1, b1);
            c0 = fut.get();
            live_threads.fetch_sub(1);
        }
        if (!spawned) {
            c0 = build(first, nl, depth + 1, b0);
            c1 = build(mid, nr, depth + 1, b1);
        }
        BvhNode64 &nd = nodes[me];
        for (int k = 0; k < 3; ++k) {
            nd.lo0[k] = b0.lo[k];
            nd.hi0[k] = b0.hi[k];
            nd.lo1[k] = b1.lo[k];
            nd.hi1[k] = b1.hi[k];
        }
        nd.c0 = c0;
        nd.c1 = c1;
        nd.pad0 = nd.pad1 = 0;
        ranges[2 * (size_t)me] = {first, nl};
        ranges[2 * (size_t)me + 1] = {mid, nr};
        return (int32_t)me;
    }
};

} // namespace

void build_bvh(const float *pos9, uint64_t n, int leaf_max, BvhResult &out) {
    out = BvhResult();
    for (int k = 0; k < 3; ++k) {
        out.lo[k] = FLT_MAX;
        out.hi[k] = -FLT_MAX;
    }
    if (n == 0) {
        out.root = 0;
        return;
    }
    Builder B;
    B.pos9 = pos9;
    B.n = n;
    B.leaf_max = std::max(1, std::min(leaf_max, 16));
    // Measured on B200 (profiles/r2_ab7_bvh_bins.txt): on the 524 k-triangle stand-in 64 bins instead of 16 are worth +5 % full-path
    // throughput over the BVH and +10 % over the Subgrids derived from it (saturated: 128 / 256 bins add nothing); the 202-triangle
    // Cornell box prefers 16.  PTB_BVH_BINS overrides.
    B.bins = n >= (1u << 15) ? 64 : 16;
    if (const char *e = getenv("PTB_BVH_BINS")) B.bins = std::max(4, std::min(256, atoi(e)));
    B.hw_threads = (int)std::max(1u, std::thread::hardware_concurrency());
    B.tbox.resize(n);
    B.cent.resize(3 * n);
    B.idx.resize(n);
    for (uint64_t i = 0; i < n; ++i) {
        const float *p = pos9 + 9 * i;
        Box b;
        b.grow(p);
        b.grow(p + 3);
        b.grow(p + 6);
        B.tbox[i] = b;
        for (int k = 0; k < 3; ++k) B.cent[3 * i + k] = 0.5f * (b.lo[k] + b.hi[k]);
        B.idx[i] = (uint32_t)i;
    }
    B.nodes.resize(n > 1 ? n - 1 : 1);
    B.ranges.resize(2 * B.nodes.size());
    Box rootBox;
    out.root = B.build(0, (uint32_t)n, 0, rootBox);
    const uint32_t used = B.next.load();
    B.nodes.resize(used);
    B.ranges.resize(2 * (size_t)used);
    // Depth-first relayout.  Indices are handed out before a node's children are built, so a single thread numbers in pre-order
    // already; but the top levels of a large scene build their subtrees concurrently (count > 200000) and those subtrees'
    // allocations interleave node by node.  Renumbering in strict pre-order puts a node's first child in the next 64 bytes —
    // the same 128-byte line every other time — for a walk that waits on its dependent node fetch (PTB_BVH_RELAYOUT=0: off).
    const char *rl = getenv("PTB_BVH_RELAYOUT");
    if (out.root >= 0 && used > 1 && !(rl && atoi(rl) == 0)) {
        std::vector<int32_t> newIdx(used, -1), order;
        order.reserve(used);
        std::vector<int32_t> todo{out.root};
        while (!todo.empty()) {
            const int32_t i = todo.back();
            todo.pop_back();
            newIdx[i] = (int32_t)order.size();
            order.push_back(i);
            const BvhNode64 &nd = B.nodes[i];
            if (nd.c1 >= 0) todo.push_back(nd.c1); // popped second: the right subtree follows the whole left one
            if (nd.c0 >= 0) todo.push_back(nd.c0);
        }
        std::vector<BvhNode64> nn(order.size());
        std::vector<BvhRange> rr(2 * order.size());
        for (size_t k = 0; k < order.size(); ++k) {
            BvhNode64 nd = B.nodes[order[k]];
            if (nd.c0 >= 0) nd.c0 = newIdx[nd.c0];
            if (nd.c1 >= 0) nd.c1 = newIdx[nd.c1];
            nn[k] = nd;
            rr[2 * k] = B.ranges[2 * (size_t)order[k]];
            rr[2 * k + 1] = B.ranges[2 * (size_t)order[k] + 1];
        }
        out.root = 0;
        B.nodes.swap(nn);
        B.ranges.swap(rr);
    }
    out.nodes.swap(B.nodes);
    out.ranges.swap(B.ranges);
    out.order.swap(B.idx);
    out.max_depth = B.max_depth.load();
    for (int k = 0; k < 3; ++k) {
        out.lo[k] = rootBox.lo[k];
        out.hi[k] = rootBox.hi[k];
    }
}

} // namespace ptb
