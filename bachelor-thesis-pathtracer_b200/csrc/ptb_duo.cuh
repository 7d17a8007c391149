// ptb_duo.cuh — the persistent-lane path kernel with TWO paths per lane.
//
// k_mega (ptb_mega.cuh) keeps one path per lane; ncu shows what that costs (profiles/r1_ncu_mega_*.txt): a lane
// whose ray is finished waits for the warp's next logic pass, and the logic pass waits for 26 such lanes, so the
// traversal code runs at 12 of 32 lanes on average, raygen at 8.6 and the rare material branches at 3.
//
// Here every lane owns two path slots (state in thread-local memory, i.e. L1) and walks at most one of them in
// registers.  A lane whose ray is finished parks the hit in the slot and goes on with its other path, so
// waiting for a batch costs nothing, and the three kinds of work are each run for a whole batch of lanes:
//   sample pass  — claim a work item (one atomicAdd per warp), check that the pixel's previous sample has landed,
//                  raygen (tracer.cpp:72-88);
//   logic pass   — shade / bounce / roulette / accumulate for one finished ray per lane (tracer.cpp:113-160);
//   walk         — while-while BVH rounds (BvhWalk::round) for the lanes that hold a ray.
// A pass runs when `keep` lanes have work of its kind, or when fewer than `walk_min` lanes can walk and at least
// `refill_min` lanes wait for a pass.  The arithmetic is the shared code of ptb_kernels.cuh / ptb_device.cuh, and
// the samples of a pixel stay serial (same readiness protocol as k_mega), so frames are bit-identical.
#pragma once
#include "ptb_mega.cuh"

#ifndef PTB_DUO_MINBLOCKS
#define PTB_DUO_MINBLOCKS 10
#endif

namespace ptb {

enum : int {
    SL_EMPTY = 0,   // no work item (a new one is claimed while the cursor has any)
    SL_WAIT = 1,    // item claimed, the pixel's previous chunk has not landed yet
    SL_NEXT = 2,    // needs raygen
    SL_READY = 3,   // path ray prepared in the slot
    SL_RSHADOW = 4, // shadow ray prepared in the slot
    SL_WALK = 5,    // path ray in the lane's registers
    SL_WSHADOW = 6, // shadow ray in the lane's registers
    SL_LOGIC = 7    // hit parked in the slot, waits for the logic pass
};

struct DuoSlotBase {
    PathState ps;
    f3 inv; // 1/direction of the prepared path ray: divided in the batched passes, not at walk set-up
    uint32_t pixel, item;
    int samplesLeft;
    float ht, hu, hv;
    int htri;
    uint32_t nTri, nAS; // tests of the parked ray (COUNT only)
};
template <bool NEE>
struct DuoSlot : DuoSlotBase {};
template <>
struct DuoSlot<true> : DuoSlotBase {
    ShadowRec sr;
    f3 sinv; // 1/direction of the shadow ray
    float shadowT;
    int shadowTri;
};

template <int MODE, bool COUNT, bool NEE>
__global__ void __launch_bounds__(128, PTB_DUO_MINBLOCKS) k_duo(ParamsDev P, SceneDev S, AttrDev A, FrameDev F, MegaArgs M) {
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    DuoSlot<NEE> sl[2];
    int st0 = SL_EMPTY, st1 = SL_EMPTY; // slot states live in registers: the scheduler reads them every turn
    int cur = -1;                       // slot whose ray is in `w`
    bool more = true;                   // warp-uniform: the cursor still has items
    BvhWalk w;
    w.ref = PTB_REF_DONE;
    int2 stackMem[PTB_STACK];
    const LocalStack stack{stackMem};
    uint32_t nTri = 0, nAS = 0;
    uint32_t cPaths = 0, cRef = 0, cRays = 0, cShadow = 0;
    unsigned long long cTri = 0, cAS = 0;

    auto state = [&](int k) { return k ? st1 : st0; };
    auto set_state = [&](int k, int v) {
        if (k) st1 = v;
        else st0 = v;
    };
    // an item's pixel is ready once its previous chunk is in HBM (k_mega's protocol)
    auto check_ready = [&](int k) {
        const uint32_t pixel = sl[k].pixel;
        const int chunk = (int)(sl[k].item / F.n_slots);
        const int s = __ldcg(&F.samples[pixel]);
        const int expect = M.first_sample + chunk * (int)M.chunk_samples;
        if (!M.check_ready || s >= expect || s >= P.target_samples) {
            const int left = min((int)M.chunk_samples, P.target_samples - s);
            if (left <= 0) {
                set_state(k, SL_EMPTY);
            } else {
                sl[k].samplesLeft = left;
                sl[k].ps.rng = __ldcg(&F.rng[pixel]);
                set_state(k, SL_NEXT);
            }
        }
    };

    for (;;) {
        const bool canWalk = cur >= 0 || st0 == SL_READY || st0 == SL_RSHADOW || st1 == SL_READY || st1 == SL_RSHADOW;
        const bool hasLogic = st0 == SL_LOGIC || st1 == SL_LOGIC;
        const bool hasSample = st0 == SL_NEXT || st1 == SL_NEXT || (more && (st0 == SL_EMPTY || st1 == SL_EMPTY));
        const int nW = __popc(__ballot_sync(full, canWalk));
        const int nL = __popc(__ballot_sync(full, hasLogic));
        const int nS = __popc(__ballot_sync(full, hasSample));
        int what; // 0 walk, 1 logic, 2 sample
        if (nL >= M.logic_keep) what = 1;
        else if (nS >= M.sample_keep) what = 2;
        else if (nW >= M.walk_min || (nW > 0 && nL + nS < M.refill_min)) what = 0;
        else if (nL > 0 && nL >= nS) what = 1;
        else if (nS > 0) what = 2;
        else if (nW > 0) what = 0;
        else if (__ballot_sync(full, st0 == SL_WAIT || st1 == SL_WAIT)) what = 2; // only parked items left: poll them
        else break;

        if (what == 2) {
            // ---- sample pass ------------------------------------------------------------------------
            int k = st0 == SL_NEXT ? 0 : (st1 == SL_NEXT ? 1 : -1);
            bool fetch = false;
            if (k < 0 && more) {
                k = st0 == SL_EMPTY ? 0 : (st1 == SL_EMPTY ? 1 : -1);
                fetch = k >= 0;
            }
            const unsigned need = __ballot_sync(full, fetch);
            if (need) {
                const int leader = __ffs(need) - 1;
                const unsigned long long want = (unsigned long long)__popc(need);
                unsigned long long base = 0;
                if (lane == leader) base = atomicAdd(M.cursor, want);
                base = __shfl_sync(full, base, leader);
                if (base + want >= M.n_items) more = false;
                if (fetch) {
                    const unsigned long long e = base + (unsigned long long)__popc(need & ((1u << lane) - 1u));
                    if (e < M.n_items) {
                        sl[k].item = (uint32_t)e;
                        sl[k].pixel = F.slot_pixel[F.slot_base + (uint32_t)e % F.n_slots];
                        set_state(k, SL_WAIT);
                    }
                }
            }
            if (st0 == SL_WAIT) check_ready(0);
            if (st1 == SL_WAIT) check_ready(1);
            if (k >= 0 && state(k) != SL_NEXT) k = st0 == SL_NEXT ? 0 : (st1 == SL_NEXT ? 1 : -1);
            if (k >= 0 && state(k) == SL_NEXT) { // raygen: setupRay + the depth-0 draw of findIntersection
                const uint32_t pixel = sl[k].pixel;
                const int x = (int)(pixel % (uint32_t)P.W), y = (int)(pixel / (uint32_t)P.W);
                PathState ps;
                ps.rng = sl[k].ps.rng;
                setup_ray(P, x, y, ps.rng, ps.o, ps.d);
                ps.tp = mk3(1.0f, 1.0f, 1.0f);
                ps.light = mk3(0.0f, 0.0f, 0.0f);
                ps.depth = 0;
                ps.flags = 0;
                roulette(ps.rng, 0, ps.tp);
                sl[k].ps = ps;
                sl[k].inv = rcp3(ps.d);
                set_state(k, SL_READY);
            }
        } else if (what == 1) {
            // ---- logic pass: one parked hit per lane ------------------------------------------------
            const int k = st0 == SL_LOGIC ? 0 : (st1 == SL_LOGIC ? 1 : -1);
            if (k >= 0) {
                PathState ps = sl[k].ps;
                const uint32_t pixel = sl[k].pixel;
                const int tri = sl[k].htri;
                bool terminated = false, shadow = false;
                if constexpr (NEE) {
                    if (ps.flags & PTB_FLAG_SHADOW) { // finish nextEventEstimation, shader.cpp:102-112 + :233 + :275
                        const ShadowRec sr = sl[k].sr;
                        const float distance = sr.distance, tAfter = sl[k].shadowT;
                        if (!(distance <= tAfter - 0.01f || distance >= tAfter + 0.01f)) {
                            const float inv_sq = smin(1.0f, (1.0f / (distance * distance)));
                            const int lm = __float_as_int(__ldg(&A.nrm[3 * sl[k].shadowTri]).w);
                            const MatDev m = A.mats[lm];
                            f3 lc = ld3(m.gamma);
                            lc = lc / 3.14f * m.emission * sr.cosSurface * inv_sq * sr.cosLight * sr.area * (float)P.n_lights;
                            ps.light = ps.light + lc * sr.tpAt;
                        }
                        ps.light = ps.light + sr.pendingEm;
                        ps.flags &= ~PTB_FLAG_SHADOW;
                    }
                }
                const bool done = (ps.flags & PTB_FLAG_DONE) != 0u;
                if (COUNT && !done) {
                    cTri += sl[k].nTri;
                    cAS += sl[k].nAS;
                }
                if (!done) cRef++;
                const bool killed = (ps.flags & PTB_FLAG_KILLED) != 0u;
                if (done) {
                    terminated = true;
                } else if (COUNT && P.debug_view) { // heat-map view, tracer.cpp:136-142
                    if (tri >= 0) {
                        ShadowOut sh;
                        f3 pendingEm;
                        shade_hit<NEE>(P, S, A, ps, sl[k].ht, tri, sl[k].hu, sl[k].hv, terminated, sh, pendingEm);
                    }
                    ps.depth++;
                    terminated = true;
                } else if (tri < 0) {
                    // miss: the reference re-traces the unchanged ray until roulette ends the path (A.3)
                    if (killed) {
                        ps.depth++;
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
                    }
                    terminated = true;
                } else {
                    ShadowOut sh;
                    f3 pendingEm;
                    shade_hit<NEE>(P, S, A, ps, sl[k].ht, tri, sl[k].hu, sl[k].hv, terminated, sh, pendingEm);
                    bool emitted = false;
                    if constexpr (NEE) {
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
                            sl[k].sr = sr;
                            sl[k].sinv = rcp3(sr.d);
                            ps.flags |= PTB_FLAG_SHADOW;
                            shadow = true;
                            emitted = true;
                        }
                    }
                    if (!emitted) ps.light = ps.light + pendingEm;
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
                        const float it = (float)(P.debug_view == 2 ? sl[k].nTri : sl[k].nAS);
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
                    const int left = sl[k].samplesLeft - 1;
                    sl[k].samplesLeft = left;
                    if (left > 0 && n < P.target_samples) {
                        F.samples[pixel] = n;
                        sl[k].ps.rng = ps.rng;
                        set_state(k, SL_NEXT);
                    } else {
                        F.rng[pixel] = ps.rng;
                        __threadfence(); // image + rng visible before the sample count releases the pixel's next chunk
                        F.samples[pixel] = n;
                        set_state(k, SL_EMPTY);
                    }
                } else {
                    if (NEE && shadow && terminated) ps.flags |= PTB_FLAG_DONE; // only the NEE term is awaited before accumulation
                    sl[k].ps = ps;
                    sl[k].inv = rcp3(ps.d);
                    set_state(k, (NEE && shadow) ? SL_RSHADOW : SL_READY);
                }
            }
        } else {
            // ---- walk, until a pass is due ---------------------------------------------------------------
            // Warp-synchronous while-while: the node loop is left as soon as `leaf_keep` lanes hold a leaf or
            // a finished ray (instead of when the last lane does), so the box tests — three quarters of the
            // traversal work — run with most lanes, and the triangle tests with the lanes that have one.
            for (;;) {
                if (cur < 0) { // take the lane's next prepared ray (the only copy of the walk set-up)
                    const int k = (st0 == SL_READY || st0 == SL_RSHADOW) ? 0 : ((st1 == SL_READY || st1 == SL_RSHADOW) ? 1 : -1);
                    if (k >= 0) {
                        bool isShadow = false;
                        if constexpr (NEE) isShadow = state(k) == SL_RSHADOW;
                        if (isShadow) {
                            if constexpr (NEE) {
                                const int lt = sl[k].sr.lightTri;
                                w.begin_inv(S, sl[k].sr.o, sl[k].sr.d, sl[k].sinv, Hit{sl[k].sr.tmax0, lt, 0.0f, 0.0f, __float_as_int(__ldg(&S.tri[3 * lt]).w)});
                            }
                            cShadow++;
                            set_state(k, SL_WSHADOW);
                        } else {
                            w.begin_inv(S, sl[k].ps.o, sl[k].ps.d, sl[k].inv, Hit{PTB_INF, -1, 0.0f, 0.0f, -1});
                            nTri = nAS = 0;
                            cRays++;
                            set_state(k, SL_WALK);
                        }
                        cur = k;
                    }
                }
                const bool isMain = cur >= 0 && state(cur) == SL_WALK;
                if (MODE == 1) {
                    for (;;) {
                        const bool inNode = cur >= 0 && w.ref >= 0;
                        if (!__ballot_sync(full, inNode)) break;
                        if (__popc(__ballot_sync(full, cur >= 0 && w.ref < 0)) >= M.leaf_keep) break;
                        if (inNode) {
                            if (COUNT && isMain) w.node_step<true>(S, stack, nAS);
                            else w.node_step<false>(S, stack, nAS);
                        }
                    }
                    if (cur >= 0 && w.ref < 0 && w.ref != PTB_REF_DONE) {
                        if (COUNT && isMain) w.leaf<true>(S, stack, nTri);
                        else w.leaf<false>(S, stack, nTri);
                    }
                } else if (cur >= 0) {
                    if (COUNT && isMain) closest_hit<MODE, true>(S, w.o, w.d, w.h, nTri, nAS);
                    else {
                        uint32_t a = 0, b = 0;
                        closest_hit<MODE, false>(S, w.o, w.d, w.h, a, b);
                    }
                    w.ref = PTB_REF_DONE;
                }
                if (cur >= 0 && w.ref == PTB_REF_DONE) {
                    bool park = true;
                    if constexpr (NEE) {
                        if (!isMain) {
                            sl[cur].shadowT = w.h.t;
                            sl[cur].shadowTri = w.h.tri;
                            park = false;
                            // the DONE branch of the logic pass ignores the path hit; otherwise the path ray is next
                            set_state(cur, (sl[cur].ps.flags & PTB_FLAG_DONE) ? SL_LOGIC : SL_READY);
                        }
                    }
                    if (park) {
                        sl[cur].ht = w.h.t;
                        sl[cur].htri = w.h.tri;
                        sl[cur].hu = w.h.u;
                        sl[cur].hv = w.h.v;
                        if (COUNT) {
                            sl[cur].nTri = nTri;
                            sl[cur].nAS = nAS;
                        }
                        set_state(cur, SL_LOGIC);
                    }
                    cur = -1;
                }
                const int walkers = __popc(__ballot_sync(full, cur >= 0 || st0 == SL_READY || st0 == SL_RSHADOW || st1 == SL_READY || st1 == SL_RSHADOW));
                const int waitL = __popc(__ballot_sync(full, st0 == SL_LOGIC || st1 == SL_LOGIC));
                if (walkers == 0 || waitL >= M.logic_keep || (walkers < M.walk_min && waitL + nS >= M.refill_min)) break;
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
