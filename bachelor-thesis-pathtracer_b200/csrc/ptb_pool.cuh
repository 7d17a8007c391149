// ptb_pool.cuh — warp-local wavefront: the persistent path kernel, second generation.
//
// k_mega (ptb_mega.cuh) removed the grid-wide barrier per bounce but ties one path to one lane, so
// a warp's traversal phase runs from 32 walking lanes down to the few longest rays before the next
// logic pass: ncu showed 12.3 (Cornell) and 8.9 (Lucy stand-in) active lanes per instruction
// (profiles/r1_ncu_mega_*.txt).  Here a warp owns a POOL of NP > 32 paths whose state lives in
// shared memory; lanes are workers:
//
//   * a lane whose ray finished posts the hit to the pool and immediately takes another path whose
//     ray is ready (warp ballot + popc compaction into a shared list) — traversal stays wide;
//   * a logic pass runs when `logic_keep` paths of the pool wait for it (or nothing else can make
//     progress) and always handles up to 32 paths at once — shading stays wide too;
//   * free pool entries pull (pixel, sample) work items from the global cursor exactly as in k_mega
//     (same readiness rule for the serial samples of a pixel).
// Still one launch per render, no grid-wide barrier, no HBM traffic for path state.
// Arithmetic = the shared device functions (setup_ray, roulette, shade_hit, BvhWalk, closest_hit):
// results are bit-identical to k_mega, the wavefront path and the oracle.
#pragma once
#include "ptb_mega.cuh"

#ifndef PTB_POOL_MINBLOCKS
#define PTB_POOL_MINBLOCKS 8
#endif

namespace ptb {

enum : int { PS_FREE = 0, PS_WAIT = 1, PS_READY_MAIN = 2, PS_READY_SHADOW = 3, PS_WALKING = 4, PS_LOGIC = 5 };

template <int NP>
struct PoolSmem { // one per warp; SoA so that lane i touching path i is conflict free
    float o[3][NP], d[3][NP], tp[3][NP], light[3][NP];
    float ht[NP], hu[NP], hv[NP];
    int htri[NP];
    int depth[NP];
    uint32_t flags[NP], rng[NP], pixel[NP], item[NP];
    int samplesLeft[NP];
    uint32_t nTri[NP], nAS[NP];
    float shT[NP]; // shadow ray result (distance after the scene test)
    int shTri[NP];
    unsigned char state[NP];
    unsigned char list[NP];
};

// compaction of the pool entries in state `a` (or `b`) into Q.list; returns how many
template <int NP>
__device__ __forceinline__ int pool_list(PoolSmem<NP> &Q, int lane, int a, int b) {
    int n = 0;
#pragma unroll
    for (int c = 0; c < (NP + 31) / 32; ++c) {
        const int i = c * 32 + lane;
        const bool is = i < NP && (Q.state[i] == a || Q.state[i] == b);
        const unsigned m = __ballot_sync(0xffffffffu, is);
        if (is) Q.list[n + __popc(m & ((1u << lane) - 1u))] = (unsigned char)i;
        n += __popc(m);
    }
    __syncwarp();
    return n;
}
template <int NP>
__device__ __forceinline__ int pool_count(const PoolSmem<NP> &Q, int lane, int a, int b) {
    int n = 0;
#pragma unroll
    for (int c = 0; c < (NP + 31) / 32; ++c) {
        const int i = c * 32 + lane;
        n += __popc(__ballot_sync(0xffffffffu, i < NP && (Q.state[i] == a || Q.state[i] == b)));
    }
    return n;
}

template <int MODE, bool COUNT, int NP>
__global__ void __launch_bounds__(128, PTB_POOL_MINBLOCKS) k_pool(ParamsDev P, SceneDev S, AttrDev A, FrameDev F, MegaArgs M,
                                                                  ShadowRec *__restrict__ shScratch) {
    extern __shared__ __align__(16) unsigned char poolRaw[];
    PoolSmem<NP> &Q = reinterpret_cast<PoolSmem<NP> *>(poolRaw)[threadIdx.x >> 5];
    ShadowRec *SR = shScratch ? shScratch + ((size_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * NP : nullptr;
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    for (int i = lane; i < NP; i += 32) Q.state[i] = PS_FREE;
    __syncwarp();

    // the ray this lane is walking: path `cur` of the pool (or -1), `shadowWalk` = its NEE shadow ray
    int cur = -1;
    bool shadowWalk = false;
    BvhWalk w;
    int2 stackMem[PTB_STACK];
    const LocalStack stack{stackMem};
    uint32_t nTri = 0, nAS = 0;
    bool more = true;
    unsigned long long cTri = 0, cAS = 0, cPaths = 0, cRef = 0, cRays = 0, cShadow = 0;

    // raygen for path i of the pool: setupRay + the depth-0 draw of findIntersection -> READY_MAIN
    auto begin_sample = [&](int i, uint32_t rng) {
        const uint32_t pixel = Q.pixel[i];
        const int x = (int)(pixel % (uint32_t)P.W), y = (int)(pixel / (uint32_t)P.W);
        f3 o, d, tp = mk3(1.0f, 1.0f, 1.0f);
        setup_ray(P, x, y, rng, o, d);
        roulette(rng, 0, tp);
        Q.o[0][i] = o.x; Q.o[1][i] = o.y; Q.o[2][i] = o.z;
        Q.d[0][i] = d.x; Q.d[1][i] = d.y; Q.d[2][i] = d.z;
        Q.tp[0][i] = 1.0f; Q.tp[1][i] = 1.0f; Q.tp[2][i] = 1.0f;
        Q.light[0][i] = 0.0f; Q.light[1][i] = 0.0f; Q.light[2][i] = 0.0f;
        Q.depth[i] = 0;
        Q.flags[i] = 0;
        Q.rng[i] = rng;
        Q.state[i] = PS_READY_MAIN;
    };
    auto start_walk = [&](int i) {
        cur = i;
        if (Q.state[i] == PS_READY_SHADOW) {
            const ShadowRec &sr = SR[i];
            w.begin(S, sr.o, sr.d, Hit{sr.tmax0, sr.lightTri, 0.0f, 0.0f, __float_as_int(__ldg(&S.tri[3 * sr.lightTri]).w)});
            shadowWalk = true;
            cShadow++;
        } else {
            w.begin(S, mk3(Q.o[0][i], Q.o[1][i], Q.o[2][i]), mk3(Q.d[0][i], Q.d[1][i], Q.d[2][i]), Hit{PTB_INF, -1, 0.0f, 0.0f, -1});
            shadowWalk = false;
            nTri = nAS = 0;
            cRays++;
        }
        Q.state[i] = PS_WALKING;
    };

    for (;;) {
        // ---- 1. free entries pull work items ----------------------------------------------------
        if (more) {
            const int nFree = pool_count(Q, lane, PS_FREE, PS_FREE);
            if (nFree) {
                unsigned long long base = 0;
                if (lane == 0) base = atomicAdd(M.cursor, (unsigned long long)nFree);
                base = __shfl_sync(full, base, 0);
                if (base + (unsigned long long)nFree >= M.n_items) more = false;
                int off = 0;
#pragma unroll
                for (int c = 0; c < (NP + 31) / 32; ++c) {
                    const int i = c * 32 + lane;
                    const bool is = i < NP && Q.state[i] == PS_FREE;
                    const unsigned m = __ballot_sync(full, is);
                    if (is) {
                        const unsigned long long e = base + (unsigned long long)(off + __popc(m & ((1u << lane) - 1u)));
                        if (e < M.n_items) {
                            Q.item[i] = (uint32_t)e;
                            Q.pixel[i] = F.slot_pixel[F.slot_base + (uint32_t)e % F.n_slots];
                            Q.state[i] = PS_WAIT;
                        }
                    }
                    off += __popc(m);
                }
                __syncwarp();
            }
        }
        // ---- 2. parked items whose pixel is ready become paths ----------------------------------
#pragma unroll
        for (int c = 0; c < (NP + 31) / 32; ++c) {
            const int i = c * 32 + lane;
            if (i < NP && Q.state[i] == PS_WAIT) {
                const uint32_t pixel = Q.pixel[i];
                const int chunk = (int)(Q.item[i] / F.n_slots);
                const int s = __ldcg(&F.samples[pixel]);
                const int expect = M.first_sample + chunk * (int)M.chunk_samples;
                if (!M.check_ready || s >= expect || s >= P.target_samples) {
                    const int left = min((int)M.chunk_samples, P.target_samples - s);
                    if (left <= 0) {
                        Q.state[i] = PS_FREE;
                    } else {
                        Q.samplesLeft[i] = left;
                        begin_sample(i, __ldcg(&F.rng[pixel]));
                    }
                }
            }
        }
        __syncwarp();

        int nLogic = pool_count(Q, lane, PS_LOGIC, PS_LOGIC);
        int nReady = pool_count(Q, lane, PS_READY_MAIN, PS_READY_SHADOW);
        int nWalk = __popc(__ballot_sync(full, cur >= 0));
        if (!nLogic && !nReady && !nWalk) {
            if (!more && pool_count(Q, lane, PS_WAIT, PS_WAIT) == 0) break;
            continue;
        }

        if (nLogic >= M.logic_keep || (nLogic && !nReady && !nWalk)) {
            // ---- 3. logic pass over up to 32 waiting paths --------------------------------------
            const int n = pool_list(Q, lane, PS_LOGIC, PS_LOGIC);
            if (lane < n) {
                const int i = Q.list[lane];
                const uint32_t pixel = Q.pixel[i];
                PathState ps;
                ps.o = mk3(Q.o[0][i], Q.o[1][i], Q.o[2][i]);
                ps.d = mk3(Q.d[0][i], Q.d[1][i], Q.d[2][i]);
                ps.tp = mk3(Q.tp[0][i], Q.tp[1][i], Q.tp[2][i]);
                ps.light = mk3(Q.light[0][i], Q.light[1][i], Q.light[2][i]);
                ps.depth = Q.depth[i];
                ps.flags = Q.flags[i];
                ps.rng = Q.rng[i];
                const float ht = Q.ht[i], hu = Q.hu[i], hv = Q.hv[i];
                const int tri = Q.htri[i];
                const uint32_t pTri = COUNT ? Q.nTri[i] : 0u, pAS = COUNT ? Q.nAS[i] : 0u;
                bool terminated = false, shadow = false;
                if (ps.flags & PTB_FLAG_SHADOW) { // finish nextEventEstimation, shader.cpp:102-112 + :233 + :275
                    const ShadowRec &sr = SR[i];
                    const float distance = sr.distance, tAfter = Q.shT[i];
                    if (!(distance <= tAfter - 0.01f || distance >= tAfter + 0.01f)) {
                        const float inv_sq = smin(1.0f, (1.0f / (distance * distance)));
                        const int lm = __float_as_int(__ldg(&A.nrm[3 * Q.shTri[i]]).w);
                        const MatDev m = A.mats[lm];
                        f3 lc = ld3(m.gamma);
                        lc = lc / 3.14f * m.emission * sr.cosSurface * inv_sq * sr.cosLight * sr.area * (float)P.n_lights;
                        ps.light = ps.light + lc * sr.tpAt;
                    }
                    ps.light = ps.light + sr.pendingEm;
                    ps.flags &= ~PTB_FLAG_SHADOW;
                }
                const bool done = (ps.flags & PTB_FLAG_DONE) != 0u;
                if (COUNT && !done) {
                    cTri += pTri;
                    cAS += pAS;
                }
                if (!done) cRef++;
                const bool killed = (ps.flags & PTB_FLAG_KILLED) != 0u;
                if (done) {
                    terminated = true;
                } else if (COUNT && P.debug_view) { // heat-map view, tracer.cpp:136-142
                    if (tri >= 0) {
                        ShadowOut sh;
                        f3 pendingEm;
                        shade_hit(P, S, A, ps, ht, tri, hu, hv, terminated, sh, pendingEm);
                    }
                    ps.depth++;
                    terminated = true;
                } else if (tri < 0) {
                    // miss: the reference re-traces the unchanged ray until roulette ends the path (A.3)
                    if (killed) {
                        ps.depth++;
                        terminated = true;
                    } else {
                        for (;;) {
                            ps.depth++;
                            if (P.max_depth > 0 && ps.depth >= P.max_depth) break;
                            if (roulette(ps.rng, ps.depth, ps.tp)) {
                                if (MODE == 2 && P.quirks) cRef++;
                                ps.depth++;
                                break;
                            }
                            cRef++;
                        }
                        terminated = true;
                    }
                } else {
                    ShadowOut sh;
                    f3 pendingEm;
                    shade_hit(P, S, A, ps, ht, tri, hu, hv, terminated, sh, pendingEm);
                    if (sh.emit) {
                        ShadowRec sr;
                        sr.o = sh.o;
                        sr.d = sh.d;
                        sr.tmax0 = sh.tmax0;
                        sr.lightTri = sh.lightTri;
                        sr.tpAt = ps.tp;
                        sr.distance = sh.distance;
                        sr.pendingEm = pendingEm;
                        sr.cosSurface = sh.cosSurface;
                        sr.cosLight = sh.cosLight;
                        sr.area = sh.area;
                        SR[i] = sr;
                        ps.flags |= PTB_FLAG_SHADOW;
                        shadow = true;
                    } else {
                        ps.light = ps.light + pendingEm;
                    }
                    ps.depth++;
                    if (P.max_depth > 0 && ps.depth >= P.max_depth) terminated = true;
                    if (killed) terminated = true;
                }
                ps.flags &= ~PTB_FLAG_KILLED;
                if (!terminated) {
                    if (roulette(ps.rng, ps.depth, ps.tp)) {
                        if (MODE == 2 && P.quirks) ps.flags |= PTB_FLAG_KILLED; // intersectGrid has no `terminated` check (A.4)
                        else {
                            ps.depth++;
                            terminated = true;
                        }
                    }
                }
                if (terminated && !shadow) { // tracer.cpp:145-155
                    const int ns = __ldcg(&F.samples[pixel]) + 1;
                    if (COUNT && P.debug_view) {
                        const float it = (float)(P.debug_view == 2 ? pTri : pAS);
                        const f3 color = (it > (float)P.debug_scale) ? mk3(1.0f, 0.0f, 0.0f) : mk3(it, it, it) / (float)P.debug_scale;
                        F.image[pixel] = make_float4(color.x, color.y, color.z, 1.0f);
                    } else {
                        const float cs = (float)ns;
                        const float4 prev = __ldcg(&F.image[pixel]);
                        f3 color = ps.light;
                        color = color + mk3(prev.x, prev.y, prev.z) * cs;
                        color = color / (cs + 1.0f);
                        F.image[pixel] = make_float4(color.x, color.y, color.z, 1.0f);
                    }
                    cPaths++;
                    const int left = Q.samplesLeft[i] - 1;
                    Q.samplesLeft[i] = left;
                    if (left > 0 && ns < P.target_samples) {
                        F.samples[pixel] = ns;
                        begin_sample(i, ps.rng);
                    } else {
                        F.rng[pixel] = ps.rng;
                        __threadfence(); // image + rng visible before the sample count releases the pixel's next chunk
                        F.samples[pixel] = ns;
                        Q.state[i] = PS_FREE;
                    }
                } else {
                    if (shadow && terminated) ps.flags |= PTB_FLAG_DONE; // only the NEE term is awaited before accumulation
                    Q.o[0][i] = ps.o.x; Q.o[1][i] = ps.o.y; Q.o[2][i] = ps.o.z;
                    Q.d[0][i] = ps.d.x; Q.d[1][i] = ps.d.y; Q.d[2][i] = ps.d.z;
                    Q.tp[0][i] = ps.tp.x; Q.tp[1][i] = ps.tp.y; Q.tp[2][i] = ps.tp.z;
                    Q.light[0][i] = ps.light.x; Q.light[1][i] = ps.light.y; Q.light[2][i] = ps.light.z;
                    Q.depth[i] = ps.depth;
                    Q.flags[i] = ps.flags;
                    Q.rng[i] = ps.rng;
                    Q.state[i] = shadow ? PS_READY_SHADOW : PS_READY_MAIN;
                }
            }
            __syncwarp();
            continue;
        }

        // ---- 4. traversal rounds; lanes that finish take the next ready path in place ------------
        int nList = nReady ? pool_list(Q, lane, PS_READY_MAIN, PS_READY_SHADOW) : 0; // ready paths, consumed from the back
        for (;;) {
            if (nList > 0) {
                const unsigned idleMask = __ballot_sync(full, cur < 0);
                if (idleMask) {
                    if (cur < 0) {
                        const int k = __popc(idleMask & ((1u << lane) - 1u));
                        if (k < nList) start_walk(Q.list[nList - 1 - k]);
                    }
                    const int took = min(nList, __popc(idleMask));
                    nList -= took;
                    nWalk += took;
                }
            }
            bool finished = false;
            if (cur >= 0) {
                bool cont;
                if (MODE == 1) {
                    if (COUNT && !shadowWalk) cont = w.round<true>(S, stack, nTri, nAS);
                    else cont = w.round<false>(S, stack, nTri, nAS);
                } else {
                    if (COUNT && !shadowWalk) closest_hit<MODE, true>(S, w.o, w.d, w.h, nTri, nAS);
                    else {
                        uint32_t a = 0, b = 0;
                        closest_hit<MODE, false>(S, w.o, w.d, w.h, a, b);
                    }
                    cont = false;
                }
                if (!cont) {
                    const int i = cur;
                    if (shadowWalk) {
                        Q.shT[i] = w.h.t;
                        Q.shTri[i] = w.h.tri;
                        if (Q.flags[i] & PTB_FLAG_DONE) {
                            Q.state[i] = PS_LOGIC; // the DONE branch of the logic pass ignores the path hit
                            cur = -1;
                            finished = true;
                        } else { // the path ray of the same bounce follows directly
                            w.begin(S, mk3(Q.o[0][i], Q.o[1][i], Q.o[2][i]), mk3(Q.d[0][i], Q.d[1][i], Q.d[2][i]), Hit{PTB_INF, -1, 0.0f, 0.0f, -1});
                            shadowWalk = false;
                            nTri = nAS = 0;
                            cRays++;
                        }
                    } else {
                        Q.ht[i] = w.h.t;
                        Q.htri[i] = w.h.tri;
                        Q.hu[i] = w.h.u;
                        Q.hv[i] = w.h.v;
                        if (COUNT) {
                            Q.nTri[i] = nTri;
                            Q.nAS[i] = nAS;
                        }
                        Q.state[i] = PS_LOGIC;
                        cur = -1;
                        finished = true;
                    }
                }
            }
            const unsigned fin = __ballot_sync(full, finished);
            if (fin) {
                nLogic += __popc(fin);
                nWalk -= __popc(fin);
                // leave the rounds when a logic pass is due or when nobody walks and nothing is ready
                if (nLogic >= M.logic_keep || (nWalk == 0 && nList == 0)) break;
            }
        }
        __syncwarp();
    }

    for (int off = 16; off > 0; off >>= 1) {
        cTri += __shfl_down_sync(full, cTri, off);
        cAS += __shfl_down_sync(full, cAS, off);
        cPaths += __shfl_down_sync(full, cPaths, off);
        cRef += __shfl_down_sync(full, cRef, off);
        cRays += __shfl_down_sync(full, cRays, off);
        cShadow += __shfl_down_sync(full, cShadow, off);
    }
    if (lane == 0) {
        if (cPaths) atomicAdd(&M.counters[0], cPaths);
        if (cRef) atomicAdd(&M.counters[1], cRef);
        if (COUNT && cTri) atomicAdd(&M.counters[2], cTri);
        if (COUNT && cAS) atomicAdd(&M.counters[3], cAS);
        if (cRays) atomicAdd(&M.counters[5], cRays);
        if (cShadow) atomicAdd(&M.counters[6], cShadow);
    }
}

} // namespace ptb
