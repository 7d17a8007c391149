// ptb_mega.cuh — the persistent-lane path kernel: ONE launch per ptb_render.
//
// Why: the wavefront pipeline (k_extend* + k_logic per bounce) pays a grid-wide barrier per bounce.
// Measured on B200: every extend launch ends with all resident warps draining their
// last rays at a few active lanes each — a fixed 0.03 ms (Cornell) to 0.1 ms (Lucy stand-in) per
// launch, 28-35 % of the extend time at one GPU and more at 1/N of the frame per GPU — plus 80 B of
// path state written and re-read per bounce and two launches per bounce.
//
// Here every lane owns one path at a time and keeps it in registers from raygen to accumulation
// (tracer.cpp:113-160 is exactly this loop, one path per thread).  Paths never leave their lane, so
// there is no barrier, no queue and no state traffic; what remains is SIMT scheduling inside a warp:
//
//   * lanes are in one of {idle, waiting for their pixel, needs-logic, walking a shadow ray, walking
//     the path ray}; each turn of the warp loop runs EITHER one logic pass for the lanes that need it
//     OR one while-while traversal round (BvhWalk::round) for the lanes that walk;
//   * logic runs when at least `logicKeep` lanes wait for it (warp ballot + popc) or nobody walks —
//     the knob trades lane occupancy of the shading code against that of the traversal code;
//   * idle lanes pull (pixel, sample-chunk) work items from one global cursor (one atomicAdd per
//     warp refill).  Samples of a pixel are serial in the reference (the RNG stream of sample k+1
//     continues sample k's, and the running mean is order dependent), so item i = (pixel i mod n,
//     chunk i div n) may only start once the pixel's previous chunk has landed in HBM; a lane whose
//     pixel is not ready parks the item and re-checks each turn (never spins: the producer may sit
//     in the same warp).
//
// The arithmetic is the same code as the wavefront path (setup_ray, roulette, shade_hit, BvhWalk,
// closest_hit), so results are bit-identical to it and to the oracle.
#pragma once
#include "ptb_kernels.cuh"

#ifndef PTB_MEGA_MINBLOCKS
#define PTB_MEGA_MINBLOCKS 12 /* 40 registers/thread: 48 warps/SM beat fewer, fatter warps despite the spills (profiles/r1_sweep_mega.txt) */
#endif
// Per mode since round 2 (profiles/r2_ab9_regcap.txt, final tree and thresholds): the BVH walks gain 3 % on the Lucy stand-in with
// 10 resident CTAs per SM (48 registers, a third fewer spills) and nothing on Cornell; the Subgrid walk loses 3 % with them.
#ifndef PTB_MEGA_MINBLOCKS_BVH
#define PTB_MEGA_MINBLOCKS_BVH 10
#endif
#define PTB_MEGA_BLOCKS_FOR(M) (((M) == 1 || (M) == 5) ? PTB_MEGA_MINBLOCKS_BVH : PTB_MEGA_MINBLOCKS)

namespace ptb {

template <int MODE>
struct WalkOf {
    typedef BvhWalk type;
};
#ifndef PTB_GRID_ROUNDS
#define PTB_GRID_ROUNDS 1 /* k_mega<TWO_PLANE> walks the Subgrids in while-while rounds (GridWalk); 0 = one-shot per lane */
#endif
#if PTB_GRID_ROUNDS
template <>
struct WalkOf<2> {
    typedef GridWalk type;
};
#endif
#define PTB_MODE_HAS_ROUNDS(M) ((M) == 1 || (M) == 5 || ((M) == 2 && PTB_GRID_ROUNDS))
template <>
struct WalkOf<5> {
    typedef BvhWalk4 type; // mode BVH over the 4-wide collapse (ptb_device.cuh)
};

enum : int { LANE_IDLE = 0, LANE_WAIT = 1, LANE_LOGIC = 2, LANE_SHADOW = 3, LANE_MAIN = 4, LANE_NEXT = 5 /* next sample of the same pixel */ };

struct ShadowRec { // what nextEventEstimation still needs once the shadow ray is back (shader.cpp:102-112)
    f3 o, d;
    f3 tpAt, pendingEm;
    float tmax0, distance, cosSurface, cosLight, area;
    int lightTri;
};

struct MegaArgs {
    uint32_t n_items;       // n_slots * n_chunks
    uint32_t chunk_samples; // samples per work item
    int first_sample;       // sample count every owned pixel has when the render starts (uniform state)
    int check_ready;        // 1: chunked mode, wait for the pixel's previous chunk
    int logic_keep;
    int sample_keep, walk_min, refill_min, leaf_keep; // k_duo only (ptb_duo.cuh)
    unsigned long long *cursor;   // work-item cursor (zeroed per launch)
    unsigned long long *counters; // [0] paths [1] reference rays [2] tri tests [3] AS tests [5] path rays [6] shadow rays
};

// EARLY (BVH only): BvhWalk::round leaves its node loop when fewer than leaf_keep/32 of the walking lanes still descend.
// NEE = false compiles the shadow-ray state machine and its 19 words of per-lane state out (the reference's
// own --testing protocol renders with NEE off, main.cpp:101).
// TEX = true compiles the texture branches of shade_hit in (scenes that bind textures; ptb_upload_textures).
template <int MODE, bool COUNT, bool NEE, bool EARLY = false, bool TEX = false>
__global__ void __launch_bounds__(128, PTB_MEGA_BLOCKS_FOR(MODE)) k_mega(ParamsDev P, SceneDev S, AttrDev A, FrameDev F, MegaArgs M) {
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    int phase = LANE_IDLE;
    bool more = true;
    uint32_t pixel = 0, item = 0;
    int samplesLeft = 0;
    PathState ps;
    ps.flags = 0;
    ps.depth = 0;
    ps.rng = 0;
    ShadowRec sr;
    float shadowT = 0.0f;
    int shadowTri = -1;
    float shadowU = 0.0f, shadowV = 0.0f; // TEX only: where the shadow ray ended (a textured light is sampled there)
    typename WalkOf<MODE>::type w;
#if PTB_SSTACK > 0
    // the first PTB_SSTACK stack entries of every lane live in shared memory (BVH mode; the other modes keep their own stacks)
    __shared__ int2 sstack[(MODE == 1 || MODE == 5) ? PTB_SSTACK * PTB_CTA : 1];
    int2 stackMem[PTB_STACK - PTB_SSTACK];
    const HybridStack<PTB_SSTACK> stack{sstack + ((MODE == 1 || MODE == 5) ? threadIdx.x : 0), stackMem};
#else
    int2 stackMem[PTB_STACK];
    const LocalStack stack{stackMem};
#endif
    uint32_t nTri = 0, nAS = 0;
    uint32_t cPaths = 0, cRef = 0, cRays = 0, cShadow = 0; // per lane and launch: far below 2^32
    unsigned long long cTri = 0, cAS = 0;

    // begin the traversal of the lane's current ray
    auto begin_main = [&]() {
        w.begin(S, ps.o, ps.d, Hit{PTB_INF, -1, 0.0f, 0.0f, -1});
        nTri = nAS = 0;
        cRays++;
        phase = LANE_MAIN;
    };
    auto begin_shadow = [&]() {
        w.begin(S, sr.o, sr.d, Hit{sr.tmax0, sr.lightTri, 0.0f, 0.0f, __float_as_int(__ldg(&S.tri[3 * sr.lightTri]).w)});
        cShadow++;
        phase = LANE_SHADOW;
    };
    // next sample of the lane's pixel (setupRay + the depth-0 draw of findIntersection)
    auto begin_sample = [&]() {
        const int x = (int)(pixel % (uint32_t)P.W), y = (int)(pixel / (uint32_t)P.W);
        setup_ray(P, x, y, ps.rng, ps.o, ps.d);
        ps.tp = mk3(1.0f, 1.0f, 1.0f);
        ps.light = mk3(0.0f, 0.0f, 0.0f);
        ps.depth = 0;
        ps.flags = 0;
        roulette(ps.rng, 0, ps.tp);
        begin_main();
    };

    for (;;) {
        // ---- A. work items for idle lanes -------------------------------------------------------
        const unsigned idle = __ballot_sync(full, phase == LANE_IDLE);
        if (more && idle) {
            const int leader = __ffs(idle) - 1;
            const unsigned long long want = (unsigned long long)__popc(idle);
            unsigned long long base = 0;
            if (lane == leader) base = atomicAdd(M.cursor, want);
            base = __shfl_sync(full, base, leader);
            if (base + want >= M.n_items) more = false;
            if (phase == LANE_IDLE) {
                const unsigned long long e = base + (unsigned long long)__popc(idle & ((1u << lane) - 1u));
                if (e < M.n_items) {
                    item = (uint32_t)e;
                    pixel = F.slot_pixel[F.slot_base + item % F.n_slots];
                    phase = LANE_WAIT;
                }
            }
        }
        if (phase == LANE_WAIT) {
            const int chunk = (int)(item / F.n_slots);
            const int s = __ldcg(&F.samples[pixel]);
            const int expect = M.first_sample + chunk * (int)M.chunk_samples;
            if (!M.check_ready || s >= expect || s >= P.target_samples) {
                samplesLeft = min((int)M.chunk_samples, P.target_samples - s);
                if (samplesLeft <= 0) {
                    phase = LANE_IDLE;
                } else {
                    ps.rng = __ldcg(&F.rng[pixel]);
                    phase = LANE_NEXT;
                }
            }
        }
        if (phase == LANE_NEXT) begin_sample(); // the only copy of the raygen code in the kernel (I-cache)

        const unsigned needLogic = __ballot_sync(full, phase == LANE_LOGIC);
        const unsigned walking = __ballot_sync(full, phase == LANE_SHADOW || phase == LANE_MAIN);
        if (!needLogic && !walking) {
            if (!more && !__ballot_sync(full, phase == LANE_WAIT)) break;
            continue;
        }

        if (needLogic && (!walking || __popc(needLogic) >= M.logic_keep)) {
            // ---- B. logic pass: the body of k_logic on register state ---------------------------
            if (phase == LANE_LOGIC) {
                bool terminated = false, shadow = false;
                if (NEE && (ps.flags & PTB_FLAG_SHADOW)) { // finish nextEventEstimation, shader.cpp:102-112 + :233 + :275
                    const float distance = sr.distance, tAfter = shadowT;
                    if (!(distance <= tAfter - 0.01f || distance >= tAfter + 0.01f)) {
                        const float inv_sq = smin(1.0f, (1.0f / (distance * distance)));
                        const int lm = __float_as_int(__ldg(&A.nrm[3 * shadowTri]).w);
                        const MatDev m = A.mats[lm];
                        f3 lc = ld3(m.gamma);
                        if (TEX && m.tex >= 0) { // getColorOfMaterial(shadowRay, lightMaterial), shader.cpp:106
                            float su = shadowU, sv = shadowV;
                            if (shadowTri == sr.lightTri) { // the light triangle's own hit was the ray's starting hit: (u,v) as at shader.cpp:97
                                Hit lh{PTB_INF, -1, 0.0f, 0.0f, -1};
                                uint32_t dummy = 0;
                                intersect_tri<false>(S.tri, sr.lightTri, sr.o, sr.d, lh, dummy);
                                su = lh.u;
                                sv = lh.v;
                            }
                            const float4 fg = tex_at(A, m.tex, hit_uv(A, shadowTri, su, sv));
                            uint32_t rs = 0; // the shadow ray's own stream: one draw in findIntersection, then the alpha draw
                            fast_random(rs);
                            const float xa = fast_random(rs);
                            // a see-through texel returns {-1,0,0}, which gammaCorrect turns into (NaN,0,0) BEFORE the
                            // reference tests for -1 (shader.cpp:107-109): the NaN is carried on, here too
                            lc = (xa < 1 - fg.w) ? mk3(__int_as_float(0x7fc00000), 0.0f, 0.0f) : mk3(fg.x, fg.y, fg.z);
                        }
                        lc = lc / 3.14f * m.emission * sr.cosSurface * inv_sq * sr.cosLight * sr.area * (float)P.n_lights;
                        ps.light = ps.light + lc * sr.tpAt;
                    }
                    ps.light = ps.light + sr.pendingEm;
                    ps.flags &= ~PTB_FLAG_SHADOW;
                }
                const int tri = w.h.tri;
                const bool done = (ps.flags & PTB_FLAG_DONE) != 0u;
                if (COUNT && !done) {
                    cTri += nTri;
                    cAS += nAS;
                }
                if (!done) cRef++;
                const bool killed = (ps.flags & PTB_FLAG_KILLED) != 0u;
                if (done) {
                    terminated = true;
                } else if (COUNT && P.debug_view) { // heat-map view, tracer.cpp:136-142
                    if (tri >= 0) {
                        ShadowOut sh;
                        f3 pendingEm;
                        shade_hit<NEE, TEX>(P, S, A, ps, w.h.t, tri, w.h.u, w.h.v, terminated, sh, pendingEm);
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
                            // the reference walks its structure again for the unchanged ray: same tests (counted when the structure
                            // is the reference's own: its tree, MODE 4, or its grids, ptb_upload_subgrid)
                            const bool recount = COUNT && (MODE == 4 || (MODE == 2 && S.ref_grids));
                            if (roulette(ps.rng, ps.depth, ps.tp)) {
                                if (MODE == 2 && P.quirks) { // TWO_PLANE traverses the killed ray once more (A.4)
                                    cRef++;
                                    if (recount) {
                                        cTri += nTri;
                                        cAS += nAS;
                                    }
                                }
                                ps.depth++;
                                break;
                            }
                            cRef++;
                            if (recount) {
                                cTri += nTri;
                                cAS += nAS;
                            }
                        }
                        terminated = true;
                    }
                } else {
                    ShadowOut sh;
                    f3 pendingEm;
                    shade_hit<NEE, TEX>(P, S, A, ps, w.h.t, tri, w.h.u, w.h.v, terminated, sh, pendingEm);
                    if (NEE && sh.emit) {
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
                    const int n = __ldcg(&F.samples[pixel]) + 1;
                    if (COUNT && P.debug_view) {
                        const float it = (float)(P.debug_view == 2 ? nTri : nAS);
                        const f3 color = (it > (float)P.debug_scale) ? mk3(1.0f, 0.0f, 0.0f) : mk3(it, it, it) / (float)P.debug_scale;
                        F.image[pixel] = make_float4(color.x, color.y, color.z, 1.0f);
                    } else {
                        const float cs = (float)n;
                        const float4 prev = __ldcg(&F.image[pixel]);
                        f3 color = ps.light;
                        color = color + mk3(prev.x, prev.y, prev.z) * cs;
                        color = color / (cs + 1.0f);
                        F.image[pixel] = make_float4(color.x, color.y, color.z, 1.0f);
                    }
                    cPaths++;
                    samplesLeft--;
                    if (samplesLeft > 0 && n < P.target_samples) {
                        F.samples[pixel] = n;
                        phase = LANE_NEXT;
                    } else {
                        F.rng[pixel] = ps.rng;
                        __threadfence(); // image + rng visible before the sample count releases the pixel's next chunk
                        F.samples[pixel] = n;
                        phase = LANE_IDLE;
                    }
                } else if (NEE && shadow) {
                    if (terminated) ps.flags |= PTB_FLAG_DONE; // only the NEE term is awaited before accumulation
                    begin_shadow();
                } else {
                    begin_main();
                }
            }
        } else {
            // ---- C. traversal rounds for the walking lanes, until enough lanes wait for logic ------
            int minActive = EARLY ? (__popc(walking) * M.leaf_keep) >> 5 : 0; // fewer descending lanes than this: go test the leaves
            for (;;) {
                if (phase == LANE_SHADOW || phase == LANE_MAIN) {
                    bool cont;
                    if (PTB_MODE_HAS_ROUNDS(MODE)) {
                        if (COUNT && phase == LANE_MAIN) cont = w.template round<true, EARLY>(S, stack, nTri, nAS, minActive);
                        else cont = w.template round<false, EARLY>(S, stack, nTri, nAS, minActive);
                    } else {
                        if (COUNT && phase == LANE_MAIN) closest_hit<MODE, true>(S, w.o, w.d, w.h, nTri, nAS);
                        else {
                            uint32_t a = 0, b = 0;
                            closest_hit<MODE, false>(S, w.o, w.d, w.h, a, b);
                        }
                        cont = false;
                    }
                    if (!cont) {
                        if (NEE && phase == LANE_SHADOW) {
                            shadowT = w.h.t;
                            shadowTri = w.h.tri;
                            if (TEX) {
                                shadowU = w.h.u;
                                shadowV = w.h.v;
                            }
                            if (ps.flags & PTB_FLAG_DONE) phase = LANE_LOGIC; // the DONE branch of the logic pass ignores the path hit
                            else begin_main();
                        } else {
                            phase = LANE_LOGIC;
                        }
                    }
                }
                const unsigned stillWalking = __ballot_sync(full, phase == LANE_SHADOW || phase == LANE_MAIN);
                const unsigned waiting = __ballot_sync(full, phase == LANE_LOGIC);
                if (!stillWalking || __popc(waiting) >= M.logic_keep) break;
                if (EARLY) minActive = (__popc(stillWalking) * M.leaf_keep) >> 5;
            }
        }
    }

    unsigned long long tPaths = cPaths, tRef = cRef, tRays = cRays, tShadow = cShadow;
    for (int off = 16; off > 0; off >>= 1) {
        cTri += __shfl_down_sync(full, cTri, off);
        cAS += __shfl_down_sync(full, cAS, off);
        tPaths += __shfl_down_sync(full, tPaths, off);
        tRef += __shfl_down_sync(full, tRef, off);
        tRays += __shfl_down_sync(full, tRays, off);
        tShadow += __shfl_down_sync(full, tShadow, off);
    }
    if (lane == 0) {
        if (tPaths) atomicAdd(&M.counters[0], tPaths);
        if (tRef) atomicAdd(&M.counters[1], tRef);
        if (COUNT && cTri) atomicAdd(&M.counters[2], cTri);
        if (COUNT && cAS) atomicAdd(&M.counters[3], cAS);
        if (tRays) atomicAdd(&M.counters[5], tRays);
        if (tShadow) atomicAdd(&M.counters[6], tShadow);
    }
}

} // namespace ptb
