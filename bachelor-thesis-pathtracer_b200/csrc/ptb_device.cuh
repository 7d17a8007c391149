// ptb_device.cuh — device-side building blocks of the sm_100a wavefront path tracer.
//
// Arithmetic contract: every float expression that decides a hit, a
// direction or an RNG draw is evaluated in the reference's order with IEEE fp32 and NO fused
// multiply-add.  The translation unit is compiled with -fmad=false -prec-div=true
// -prec-sqrt=true -ftz=false, so `a*b+c` below is two roundings, `/` and sqrtf are correctly
// rounded, and denormals are kept — the same results as the reference's x86 SSE/AVX code
// (CMakeLists.txt:12 has no -mfma).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace ptb {

#define PTB_INF __int_as_float(0x7f800000)
#define PTB_EPS 0.00001f /* src/common.h:6 */

struct f3 {
    float x, y, z;
};
__device__ __forceinline__ f3 mk3(float x, float y, float z) { return f3{x, y, z}; }
__device__ __forceinline__ f3 operator+(f3 a, f3 b) { return f3{a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ f3 operator-(f3 a, f3 b) { return f3{a.x - b.x, a.y - b.y, a.z - b.z}; }
__device__ __forceinline__ f3 operator*(f3 a, float s) { return f3{a.x * s, a.y * s, a.z * s}; }
__device__ __forceinline__ f3 operator*(float s, f3 a) { return f3{a.x * s, a.y * s, a.z * s}; }
__device__ __forceinline__ f3 operator*(f3 a, f3 b) { return f3{a.x * b.x, a.y * b.y, a.z * b.z}; }
// x / s, bit for bit, without waking the slow path of the IEEE division for a zero numerator.  CUDA's div.rn.f32
// range check sends 0 / s to its ~100-instruction subroutine, and axis-aligned normals, unit tangents and black
// radiance put exact zeros into most normalisations and accumulations: ncu counted 4.6 M calls of that routine per
// 14 M Cornell rays, 6.7 % of all issued instructions.  (+-0) / s for s > 0 is +-0 with the numerator's sign.
// A branch around the division gets if-converted by the compiler (the division, range check and call stay
// unconditional), so the zero numerator is replaced by 1 for the division and the quotient discarded.
__device__ __forceinline__ float divz(float x, float s) {
    const bool live = x != 0.0f || !(s > 0.0f);
    float num = live ? x : 1.0f;
    asm("" : "+f"(num)); // opaque: otherwise the select is folded back into x / s
    const float q = num / s;
    return live ? q : x;
}
__device__ __forceinline__ f3 operator/(f3 a, float s) { return f3{divz(a.x, s), divz(a.y, s), divz(a.z, s)}; }
// Eigen Vector3f::dot order: a0*b0 + (a1*b1 + a2*b2)   
__device__ __forceinline__ float dot3(f3 a, f3 b) { return a.x * b.x + (a.y * b.y + a.z * b.z); }
__device__ __forceinline__ f3 cross3(f3 a, f3 b) {
    return f3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
// Eigen normalize(): squaredNorm > 0 ? v / sqrt(squaredNorm) : v   (Eigen/src/Core/Dot.h:134-138)
__device__ __forceinline__ f3 normalize3(f3 v) {
    float z = v.x * v.x + (v.y * v.y + v.z * v.z);
    if (z > 0.0f) {
        float l = sqrtf(z);
        return f3{divz(v.x, l), divz(v.y, l), divz(v.z, l)};
    }
    return v;
}
__device__ __forceinline__ float max3(f3 v) { return fmaxf(v.x, fmaxf(v.y, v.z)); } // src/types/vector.cpp:6-9
// std::max / std::min select forms (NaN-order sensitive; src/types/aabb.cpp:17-23)
__device__ __forceinline__ float smax(float a, float b) { return (a < b) ? b : a; }
__device__ __forceinline__ float smin(float a, float b) { return (b < a) ? b : a; }
__device__ __forceinline__ float comp(f3 v, int i) { return i == 0 ? v.x : (i == 1 ? v.y : v.z); }
__device__ __forceinline__ f3 rcp3(f3 d) { return f3{1.0f / d.x, 1.0f / d.y, 1.0f / d.z}; }

// ---- RNG: src/common.cpp:8-21.  (float)UINT32_MAX rounds to 2^32, so the divide is an exact
// scale by 2^-32; the result lies in [0,1] inclusive.
__device__ __forceinline__ uint32_t hash_coords(int x, int y) { return ((uint32_t)x << 16) | (uint32_t)y; }
__device__ __forceinline__ float fast_random(uint32_t &seed) {
    uint32_t state = seed * 747796405u + 2891336453u;
    uint32_t word = ((state >> ((state >> 28u) + 4u)) ^ state) * 277803737u;
    seed = (word >> 22u) ^ word;
    return __uint2float_rn(seed) * 2.3283064365386963e-10f;
}

// ---- hit record kept during traversal (deferred attributes: SURVEY.md §8 a10) -------------
struct Hit {
    float t;  // ray.tmax
    int tri;  // index into the leaf-ordered triangle array, -1 = miss
    float u, v;
    int id;   // reference triangle id of `tri` (tie rule)
};

// Flattened structures in HBM (layouts: DESIGN.md "Data layout").
struct SceneDev {
    const float4 *__restrict__ tri;     // 3 float4 per triangle: (v0, id), (e1, -), (e2, -); leaf order
    const float4 *__restrict__ nodes;   // BVH2, 4 float4 (64 B) per inner node
    const float4 *__restrict__ nodes4;  // the same tree collapsed to 4-wide nodes (k_collapse4), 8 float4 (128 B) per BVH2 node index; null = not built
    int bvh_root;                       // node ref (>=0 inner, <0 leaf)
    uint32_t n_tris;
    // Subgrid ("2Plane")
    const float4 *__restrict__ grid_hdr; // 4 float4 (64 B) per grid
    const uint2 *__restrict__ grid_lut;  // {start,end} per cell, absolute into grid_idx
    const uint32_t *__restrict__ grid_idx;
    // Subbeams
    const float4 *__restrict__ beam;     // 3 float4 (48 B) per beam: minUVST, maxUVST, (start,end,hasTris,-)
    const uint32_t *__restrict__ beam_root[3];
    const uint32_t *__restrict__ beam_idx;
    float beam_gmin[3][3], beam_gmax[3][3], beam_amin[3], beam_amax[3];
    uint32_t beam_grid_size;
    uint32_t n_beams;
    // the reference's own two-level BVH (ptb_upload_ref_bvh), traversed node for node like bvh.cpp:48-99 so that
    // the two test counters and the heat maps are the reference's numbers
    const float4 *__restrict__ ref_node;     // 3 float4 (48 B) per node: (bmin, start) (bmax, end) (left, right, hasTris, -)
    const uint32_t *__restrict__ ref_idx;    // bvh.cpp:20 `indicies`: triangle ids (hasTris) or object ids
    const int *__restrict__ ref_obj_root;    // Object::root per object
    const uint32_t *__restrict__ ref_tri_pos; // reference triangle id -> leaf position in `tri`
    int ref_root;
    int ref_grids; // the Subgrid pools hold the reference's own hierarchy (ptb_upload_subgrid): count re-traced misses like the CPU
};

// ---- Moeller-Trumbore: src/primitives/triangle.cpp:7-52, positions only --------------------
// Accept rule: !(t < 1e-5) && !(tmax < t); among bit-equal t the larger reference id wins
// (the reference lets the later-tested one win; brute-force order == ascending id).
//
// The reference divides first (inv = 1/det) and then rejects on u, v.  Most tests fail on u or v,
// so the sign/magnitude of the un-divided numerators is checked first and the IEEE division is
// only paid by candidates that can still hit.  The pre-rejects are conservative: they fire only
// where the exactly rounded u (or v) is certainly outside [0,1]:
//   du*det < -1e-20  =>  du, det nonzero with opposite signs and |du/det| far above the smallest
//                        denormal for any |det| < 3e12, so RN(du*RN(1/det)) < 0;
//   |du| > |det|*(1+1e-6)  =>  RN(du*RN(1/det)) > 1  (two roundings, 2^-23 relative in total).
// Everything that survives is evaluated in the reference's order, bit for bit.
template <bool COUNT>
__device__ __forceinline__ void intersect_tri(const float4 *__restrict__ tri, int idx, f3 o, f3 d, Hit &h, uint32_t &nTri) {
    const float4 a = __ldg(&tri[3 * idx]);
    const float4 b = __ldg(&tri[3 * idx + 1]);
    const float4 c = __ldg(&tri[3 * idx + 2]);
    if (COUNT) nTri++;
    const f3 v0 = mk3(a.x, a.y, a.z), e1 = mk3(b.x, b.y, b.z), e2 = mk3(c.x, c.y, c.z);
    const f3 p = cross3(d, e2);
    const float det = dot3(e1, p);
    const float adet = fabsf(det);
    if (adet <= 0.01f) return; // fabs(det) < 0.01 evaluated in double (triangle.cpp:20)
    const f3 tv = o - v0;
    const float du = dot3(tv, p);
    if (du * det < -1e-20f || fabsf(du) > adet * 1.000001f) return;
    const f3 q = cross3(tv, e1);
    const float dv = dot3(d, q);
    if (dv * det < -1e-20f || fabsf(dv) > adet * 1.000001f) return;
    const float inv = 1.0f / det;
    const float u = du * inv;
    if (0.0f > u || u > 1.0f) return;
    const float v = dv * inv;
    if (0.0f > v || u + v > 1.0f) return;
    const float t = dot3(e2, q) * inv;
    if (t < 0.00001f || h.t < t) return;
    const int id = __float_as_int(a.w);
    if (t == h.t && id < h.id) return;
    h.t = t;
    h.tri = idx;
    h.u = u;
    h.v = v;
    h.id = id;
}

// ---- slab test: src/types/aabb.cpp:9-45 (same nesting of select-min/max) --------------------
__device__ __forceinline__ bool slab(f3 lo, f3 hi, f3 o, f3 inv, float tmax, float &tn) {
    const float tx1 = (lo.x - o.x) * inv.x, tx2 = (hi.x - o.x) * inv.x;
    const float ty1 = (lo.y - o.y) * inv.y, ty2 = (hi.y - o.y) * inv.y;
    const float tz1 = (lo.z - o.z) * inv.z, tz2 = (hi.z - o.z) * inv.z;
    const float t0 = smax(smax(smax(smin(tx1, tx2), smin(ty1, ty2)), smin(tz1, tz2)), 0.0f);
    const float t1 = smin(smin(smin(smax(tx1, tx2), smax(ty1, ty2)), smax(tz1, tz2)), tmax);
    tn = t0;
    return t0 <= t1;
}
// Same test with FMNMX.  std::min/max (select form) and fminf/fmaxf differ only when an operand
// is NaN or in the sign of a zero result.  A NaN needs (bound - origin) == 0 times an infinite
// inv_dir component, so rays whose inv_dir is finite in all three components cannot produce one;
// the sign of a zero never changes `t0 <= t1`, the child ordering `d1 >= d0` or the pop-time cull.
__device__ __forceinline__ bool slab_fast(f3 lo, f3 hi, f3 o, f3 inv, float tmax, float &tn) {
    const float tx1 = (lo.x - o.x) * inv.x, tx2 = (hi.x - o.x) * inv.x;
    const float ty1 = (lo.y - o.y) * inv.y, ty2 = (hi.y - o.y) * inv.y;
    const float tz1 = (lo.z - o.z) * inv.z, tz2 = (hi.z - o.z) * inv.z;
    const float t0 = fmaxf(fmaxf(fmaxf(fminf(tx1, tx2), fminf(ty1, ty2)), fminf(tz1, tz2)), 0.0f);
    const float t1 = fminf(fminf(fminf(fmaxf(tx1, tx2), fmaxf(ty1, ty2)), fmaxf(tz1, tz2)), tmax);
    tn = t0;
    return t0 <= t1;
}
__device__ __forceinline__ bool finite3(f3 v) { return fabsf(v.x) < PTB_INF && fabsf(v.y) < PTB_INF && fabsf(v.z) < PTB_INF; }

// Box test for TREE NAVIGATION only, one FFMA per plane: t = bound * inv + (-(origin * inv)).  Which boxes a walk visits does not
// enter the result — the hit record comes from intersect_tri alone, in the reference's arithmetic — provided no box the
// reference-order test accepts is rejected.  The FMA form rounds origin * inv once (0.5 ulp of |origin * inv|, which can exceed
// the slab distance itself when the origin is far from the box along a near-parallel axis), so the interval is widened by
// `pad` = 2^-21 * max|origin * inv| (absolute) and 2^-22 (relative) on both ends: a superset of the exact test's acceptances.
// The entry distance only orders the children and culls at pop time; the widened (smaller) value is conservative for both.
// Callers use it only for rays whose inv_dir is finite and below 2^100 in magnitude (no inf - inf).
#ifndef PTB_SLAB_FMA
#define PTB_SLAB_FMA 0
#endif
__device__ __forceinline__ bool slab_fma(f3 lo, f3 hi, f3 inv, f3 noi, float pad, float tmax, float &tn) {
    const float tx1 = __fmaf_rn(lo.x, inv.x, noi.x), tx2 = __fmaf_rn(hi.x, inv.x, noi.x);
    const float ty1 = __fmaf_rn(lo.y, inv.y, noi.y), ty2 = __fmaf_rn(hi.y, inv.y, noi.y);
    const float tz1 = __fmaf_rn(lo.z, inv.z, noi.z), tz2 = __fmaf_rn(hi.z, inv.z, noi.z);
    const float a = fmaxf(fmaxf(fminf(tx1, tx2), fminf(ty1, ty2)), fminf(tz1, tz2));
    const float b = fminf(fminf(fmaxf(tx1, tx2), fmaxf(ty1, ty2)), fmaxf(tz1, tz2));
    // a < 0 clamps to 0 whatever the factor; b < 0 (box behind the origin) only needs the absolute pad
    const float t0 = fmaxf(__fmaf_rn(a, 0.99999976f, -pad), 0.0f);
    const float t1 = fminf(__fmaf_rn(b, 1.00000024f, pad), tmax);
    tn = t0;
    return t0 <= t1;
}
__device__ __forceinline__ bool slab_fma_ok(f3 v) {
    const float lim = 1.2676506e30f; // 2^100
    return fabsf(v.x) < lim && fabsf(v.y) < lim && fabsf(v.z) < lim;
}

// ---- traversal stacks ------------------------------------------------------------------------------------------------
// LocalStack: the whole stack in local memory (per-thread, interleaved by the hardware).  HybridStack: entries [0, NS) in shared
// memory — column `tid` of an [NS][CTA threads] array, conflict-free for any mix of per-lane depths — and the overflow in local
// memory.  ncu on the round-1 kernel: 512 B of local stack per lane x 227 k lanes = 116 MB, the size of the L2, written back
// to DRAM at 1 GB per 4-spp Lucy launch and competing with node fetches for the L1 tag stage (LDL + STL 5 % of issue).
#ifndef PTB_SSTACK
#define PTB_SSTACK 0
#endif
#define PTB_CTA 128 /* threads per CTA of every traversal kernel (kThreads in ptb_api.cu) */
struct LocalStack {
    int2 *p;
    __device__ __forceinline__ int2 get(int i) const { return p[i]; }
    __device__ __forceinline__ void put(int i, int2 e) const { p[i] = e; }
};
template <int NS>
struct HybridStack {
    int2 *sm; // &shared[threadIdx.x]
    int2 *lm; // local overflow, PTB_STACK - NS entries
    __device__ __forceinline__ int2 get(int i) const { return i < NS ? sm[i * PTB_CTA] : lm[i - NS]; }
    __device__ __forceinline__ void put(int i, int2 e) const {
        if (i < NS) sm[i * PTB_CTA] = e;
        else lm[i - NS] = e;
    }
};

#ifndef PTB_STACK
#define PTB_STACK 64
#endif
#define PTB_GSTACK 160 /* Subgrid: <= 16 children pushed per level */
#define PTB_REF_DONE ((int)0x80000000) /* never a leaf reference: first < 2^27 - 1 */

// leaf reference: ~((first << 4) | (count - 1)), count <= 16
__device__ __forceinline__ void leaf_decode(int ref, int &first, int &count) {
    const uint32_t r = (uint32_t)~ref;
    first = (int)(r >> 4);
    count = (int)(r & 15u) + 1;
}

// ---- BVH closest hit: stands in for findBVHIntesection (bvh.cpp:48-99) ---------------------
// One 64-byte node holds both children's boxes, so each inner node costs 4 LDG.128 and two slab
// tests (the reference re-tests a node's own box after its parent already did).  Near child
// first; the far child is pushed with its entry distance and culled against the shrunken tmax
// when popped (same effect as the reference's pop-time re-test).
//
// The walk is a resumable state machine so that a warp can (a) run "while-while": every lane
// descends inner nodes until it holds a leaf, then all lanes test their leaves together, and
// (b) hand finished lanes new rays between rounds (k_extend's dynamic fetch).
struct BvhWalk {
    f3 o, d, inv;
    Hit h;
    int ref;   // >= 0 inner node, < 0 leaf, PTB_REF_DONE finished
    int sp;
    bool fast; // inv_dir finite: FMNMX / FMA slab test allowed
#if PTB_SLAB_FMA
    float pad; // slab_fma: 2^-21 * max|o * inv|
#endif
    // The stack itself (far child reference + its entry distance) is NOT a member: a dynamically
    // indexed array inside the struct drags every scalar above into local memory with it (seen in
    // ncu as 9 % LDL/STL).  Callers own the storage and pass a LocalStack / HybridStack to pop()/round().

    // `inv_` = the reciprocal direction, possibly computed elsewhere (k_duo prepares it in its batched passes)
    __device__ __forceinline__ void begin_inv(const SceneDev &S, f3 o_, f3 d_, f3 inv_, const Hit &h0) {
        o = o_;
        d = d_;
        inv = inv_;
#if PTB_SLAB_FMA
        fast = slab_fma_ok(inv);
        pad = fmaxf(fmaxf(fabsf(o.x * inv.x), fabsf(o.y * inv.y)), fabsf(o.z * inv.z)) * 4.7683716e-7f;
#else
        fast = finite3(inv);
#endif
        h = h0;
        sp = 0;
        ref = S.n_tris ? S.bvh_root : PTB_REF_DONE;
    }
    __device__ __forceinline__ void begin(const SceneDev &S, f3 o_, f3 d_, const Hit &h0) { begin_inv(S, o_, d_, rcp3(d_), h0); }
    template <class ST>
    __device__ __forceinline__ void pop(const ST &stack) {
        for (;;) {
            if (sp == 0) {
                ref = PTB_REF_DONE;
                return;
            }
            sp--;
            const int2 e = stack.get(sp);
            if (!(__int_as_float(e.y) > h.t)) {
                ref = e.x;
                return;
            }
        }
    }
    // one inner node (ref >= 0): both children's boxes, near child next, far child pushed
    template <bool COUNT, class ST>
    __device__ __forceinline__ void node_step(const SceneDev &S, const ST &stack, uint32_t &nBox) {
        const float4 *n = S.nodes + 4 * (size_t)ref;
        const float4 q0 = __ldg(n), q1 = __ldg(n + 1), q2 = __ldg(n + 2);
        const int4 q3 = __ldg(reinterpret_cast<const int4 *>(n + 3));
        if (COUNT) nBox += 2;
        float d0, d1;
        bool h0, h1;
        if (fast) {
#if PTB_SLAB_FMA
            const f3 noi = mk3(-(o.x * inv.x), -(o.y * inv.y), -(o.z * inv.z));
            h0 = slab_fma(mk3(q0.x, q0.y, q0.z), mk3(q0.w, q1.x, q1.y), inv, noi, pad, h.t, d0);
            h1 = slab_fma(mk3(q1.z, q1.w, q2.x), mk3(q2.y, q2.z, q2.w), inv, noi, pad, h.t, d1);
#else
            h0 = slab_fast(mk3(q0.x, q0.y, q0.z), mk3(q0.w, q1.x, q1.y), o, inv, h.t, d0);
            h1 = slab_fast(mk3(q1.z, q1.w, q2.x), mk3(q2.y, q2.z, q2.w), o, inv, h.t, d1);
#endif
        } else {
            h0 = slab(mk3(q0.x, q0.y, q0.z), mk3(q0.w, q1.x, q1.y), o, inv, h.t, d0);
            h1 = slab(mk3(q1.z, q1.w, q2.x), mk3(q2.y, q2.z, q2.w), o, inv, h.t, d1);
        }
        if (h0 && h1) {
            // bvh.cpp:75-84: left first unless the right child is strictly nearer
            const bool leftFirst = d1 >= d0;
            stack.put(sp++, make_int2(leftFirst ? q3.y : q3.x, __float_as_int(leftFirst ? d1 : d0)));
            ref = leftFirst ? q3.x : q3.y;
        } else if (h0) {
            ref = q3.x;
        } else if (h1) {
            ref = q3.y;
        } else {
            pop(stack);
        }
    }
    // one leaf (ref < 0, not PTB_REF_DONE): its triangles, then the next reference from the stack
    template <bool COUNT, class ST>
    __device__ __forceinline__ void leaf(const SceneDev &S, const ST &stack, uint32_t &nTri) {
        int first, count;
        leaf_decode(ref, first, count);
        for (int i = 0; i < count; ++i) intersect_tri<COUNT>(S.tri, first + i, o, d, h, nTri);
        pop(stack);
    }
    // one round: descend to the next leaf (or finish), test it, pop.  Returns false when done.
    // EARLY: leave the node loop once fewer than `minActive` lanes of the warp are still descending — the lanes
    // that already hold a leaf stop waiting for the last stragglers, which resume next round.  Pays on deep
    // trees (Lucy stand-in +17 %); on a 100-node tree the vote per step costs more than it saves
    // (profiles/r1_sweep_mega_leaf.txt), hence a compile-time switch.
#ifndef PTB_SPECULATE
#define PTB_SPECULATE 0
#endif
#if PTB_SPECULATE
    // Speculative variant (Aila & Laine 2009, "postponed leaf"): a lane that reaches a leaf holds it back and keeps descending
    // to a SECOND leaf before both are tested, so lanes stay in the node loop longer and the leaf phase has more work per
    // lane.  The held leaf's hit cannot cull the nodes visited meanwhile (a few more box tests); closest hits are unchanged.
    template <bool COUNT, bool EARLY = false, class ST>
    __device__ __forceinline__ bool round(const SceneDev &S, const ST &stack, uint32_t &nTri, uint32_t &nBox, int minActive = 0) {
        int held = 0; // leaf references are negative: 0 = none
        for (;;) {
            while (ref >= 0) {
                node_step<COUNT>(S, stack, nBox);
                if (EARLY && __popc(__activemask()) < minActive) break;
            }
            if (ref >= 0 || ref == PTB_REF_DONE || held != 0) break;
            held = ref;
            pop(stack);
        }
        if (held != 0) {
            int first, count;
            leaf_decode(held, first, count);
            for (int i = 0; i < count; ++i) intersect_tri<COUNT>(S.tri, first + i, o, d, h, nTri);
        }
        if (ref < 0 && ref != PTB_REF_DONE) leaf<COUNT>(S, stack, nTri);
        return ref != PTB_REF_DONE;
    }
#else
    template <bool COUNT, bool EARLY = false, class ST>
    __device__ __forceinline__ bool round(const SceneDev &S, const ST &stack, uint32_t &nTri, uint32_t &nBox, int minActive = 0) {
        while (ref >= 0) {
            node_step<COUNT>(S, stack, nBox);
            if (EARLY && __popc(__activemask()) < minActive) break;
        }
        if ((!EARLY || ref < 0) && ref != PTB_REF_DONE) leaf<COUNT>(S, stack, nTri);
        return ref != PTB_REF_DONE;
    }
#endif
};

// ---- 4-wide BVH ------------------------------------------------------------------------------------------------------
// The BVH2 collapsed by k_collapse4: node i of the wide tree = BVH2 node i with its (up to four) grandchildren as children —
// an inner child is replaced by its two children, a leaf child stays.  Same node indices as the BVH2 (only the nodes at even
// depth are ever reached), so the collapse is one independent thread per node and works for host- and GPU-built trees alike.
// Layout, 8 x float4 = 128 B (four 32-byte sectors, one cache line): lo.x[4] lo.y[4] lo.z[4] hi.x[4] hi.y[4] hi.z[4] ref[4] pad.
// Empty slots hold PTB_REF_DONE.  Why: ncu on the BVH2 kernel shows the node loop waiting on its dependent node fetch
// (stall_long_sb 47 % on the Lucy stand-in, L1 hit 68 %) and ~30 instructions of push / pop / branch per 59 of slab arithmetic;
// a 4-wide step halves the dependent fetches per ray and amortises the control over four boxes.  Closest hits do not depend on
// the visit order (ties go to the larger reference id), so frames are bit-identical to the BVH2 walk's.
// Up to three pushes per visited node: the BVH2 may be at most PTB_WIDE_MAX_DEPTH deep for PTB_STACK entries.
#define PTB_WIDE_MAX_DEPTH 40 /* 3 * ceil(40 / 2) = 60 < PTB_STACK */
struct BvhWalk4 {
    f3 o, d, inv;
    Hit h;
    int ref;
    int sp;
    bool fast;

    __device__ __forceinline__ void begin(const SceneDev &S, f3 o_, f3 d_, const Hit &h0) {
        o = o_;
        d = d_;
        inv = rcp3(d_);
        fast = finite3(inv);
        h = h0;
        sp = 0;
        ref = S.n_tris ? S.bvh_root : PTB_REF_DONE;
    }
    template <class ST>
    __device__ __forceinline__ void pop(const ST &stack) {
        for (;;) {
            if (sp == 0) {
                ref = PTB_REF_DONE;
                return;
            }
            sp--;
            const int2 e = stack.get(sp);
            if (!(__int_as_float(e.y) > h.t)) {
                ref = e.x;
                return;
            }
        }
    }
    template <bool COUNT, class ST>
    __device__ __forceinline__ void node_step(const SceneDev &S, const ST &stack, uint32_t &nBox) {
        const float4 *n = S.nodes4 + 8 * (size_t)ref;
        const float4 lx = __ldg(n), ly = __ldg(n + 1), lz = __ldg(n + 2), hx = __ldg(n + 3), hy = __ldg(n + 4), hz = __ldg(n + 5);
        const int4 c = __ldg(reinterpret_cast<const int4 *>(n + 6));
        if (COUNT) nBox += 4;
        float k0, k1, k2, k3; // entry distance, +inf where the child is missed or absent
        bool b0, b1, b2, b3;
        if (fast) {
            b0 = slab_fast(mk3(lx.x, ly.x, lz.x), mk3(hx.x, hy.x, hz.x), o, inv, h.t, k0);
            b1 = slab_fast(mk3(lx.y, ly.y, lz.y), mk3(hx.y, hy.y, hz.y), o, inv, h.t, k1);
            b2 = slab_fast(mk3(lx.z, ly.z, lz.z), mk3(hx.z, hy.z, hz.z), o, inv, h.t, k2);
            b3 = slab_fast(mk3(lx.w, ly.w, lz.w), mk3(hx.w, hy.w, hz.w), o, inv, h.t, k3);
        } else {
            b0 = slab(mk3(lx.x, ly.x, lz.x), mk3(hx.x, hy.x, hz.x), o, inv, h.t, k0);
            b1 = slab(mk3(lx.y, ly.y, lz.y), mk3(hx.y, hy.y, hz.y), o, inv, h.t, k1);
            b2 = slab(mk3(lx.z, ly.z, lz.z), mk3(hx.z, hy.z, hz.z), o, inv, h.t, k2);
            b3 = slab(mk3(lx.w, ly.w, lz.w), mk3(hx.w, hy.w, hz.w), o, inv, h.t, k3);
        }
        // slots 0 and 1 always exist (an inner BVH2 node has two children); 2 and 3 may be empty
        k0 = b0 ? k0 : PTB_INF;
        k1 = b1 ? k1 : PTB_INF;
        k2 = (b2 && c.z != PTB_REF_DONE) ? k2 : PTB_INF;
        k3 = (b3 && c.w != PTB_REF_DONE) ? k3 : PTB_INF;
        int r0 = c.x, r1 = c.y, r2 = c.z, r3 = c.w;
        // ascending by entry distance, misses (+inf) last: 5-comparator network (0,1)(2,3)(0,2)(1,3)(1,2)
#define PTB_CE(ka, ra, kb, rb)              \
    do {                                    \
        const bool sw_ = kb < ka;           \
        const float tk_ = sw_ ? kb : ka;    \
        const int tr_ = sw_ ? rb : ra;      \
        kb = sw_ ? ka : kb;                 \
        rb = sw_ ? ra : rb;                 \
        ka = tk_;                           \
        ra = tr_;                           \
    } while (0)
        PTB_CE(k0, r0, k1, r1);
        PTB_CE(k2, r2, k3, r3);
        PTB_CE(k0, r0, k2, r2);
        PTB_CE(k1, r1, k3, r3);
        PTB_CE(k1, r1, k2, r2);
#undef PTB_CE
        if (k0 == PTB_INF) { // nothing hit
            pop(stack);
            return;
        }
        // the nearest goes on, the others wait farthest-first so the nearer of them pops first
        if (k3 != PTB_INF) stack.put(sp++, make_int2(r3, __float_as_int(k3)));
        if (k2 != PTB_INF) stack.put(sp++, make_int2(r2, __float_as_int(k2)));
        if (k1 != PTB_INF) stack.put(sp++, make_int2(r1, __float_as_int(k1)));
        ref = r0;
    }
    template <bool COUNT, class ST>
    __device__ __forceinline__ void leaf(const SceneDev &S, const ST &stack, uint32_t &nTri) {
        int first, count;
        leaf_decode(ref, first, count);
        for (int i = 0; i < count; ++i) intersect_tri<COUNT>(S.tri, first + i, o, d, h, nTri);
        pop(stack);
    }
    template <bool COUNT, bool EARLY = false, class ST>
    __device__ __forceinline__ bool round(const SceneDev &S, const ST &stack, uint32_t &nTri, uint32_t &nBox, int minActive = 0) {
        while (ref >= 0) {
            node_step<COUNT>(S, stack, nBox);
            if (EARLY && __popc(__activemask()) < minActive) break;
        }
        if ((!EARLY || ref < 0) && ref != PTB_REF_DONE) leaf<COUNT>(S, stack, nTri);
        return ref != PTB_REF_DONE;
    }
};

// one thread per BVH2 node: its wide node (see BvhWalk4).  n2 = BVH2 inner nodes.
__global__ void k_collapse4(const float4 *__restrict__ nodes, float4 *__restrict__ nodes4, uint32_t n2) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += gridDim.x * blockDim.x) {
        const float4 *n = nodes + 4 * (size_t)i;
        const float4 q0 = n[0], q1 = n[1], q2 = n[2];
        const int4 q3 = *reinterpret_cast<const int4 *>(n + 3);
        float lo[4][3], hi[4][3];
        int ref[4] = {PTB_REF_DONE, PTB_REF_DONE, PTB_REF_DONE, PTB_REF_DONE};
        int k = 0;
        const float cb[2][6] = {{q0.x, q0.y, q0.z, q0.w, q1.x, q1.y}, {q1.z, q1.w, q2.x, q2.y, q2.z, q2.w}};
        const int cr[2] = {q3.x, q3.y};
        for (int s = 0; s < 2; ++s) {
            if (cr[s] >= 0) { // inner child: its two children take its place
                const float4 *m = nodes + 4 * (size_t)cr[s];
                const float4 m0 = m[0], m1 = m[1], m2 = m[2];
                const int4 m3 = *reinterpret_cast<const int4 *>(m + 3);
                lo[k][0] = m0.x; lo[k][1] = m0.y; lo[k][2] = m0.z; hi[k][0] = m0.w; hi[k][1] = m1.x; hi[k][2] = m1.y;
                ref[k++] = m3.x;
                lo[k][0] = m1.z; lo[k][1] = m1.w; lo[k][2] = m2.x; hi[k][0] = m2.y; hi[k][1] = m2.z; hi[k][2] = m2.w;
                ref[k++] = m3.y;
            } else {
                for (int a = 0; a < 3; ++a) {
                    lo[k][a] = cb[s][a];
                    hi[k][a] = cb[s][3 + a];
                }
                ref[k++] = cr[s];
            }
        }
        for (; k < 4; ++k)
            for (int a = 0; a < 3; ++a) lo[k][a] = hi[k][a] = 0.0f;
        float4 *o = nodes4 + 8 * (size_t)i;
        o[0] = make_float4(lo[0][0], lo[1][0], lo[2][0], lo[3][0]);
        o[1] = make_float4(lo[0][1], lo[1][1], lo[2][1], lo[3][1]);
        o[2] = make_float4(lo[0][2], lo[1][2], lo[2][2], lo[3][2]);
        o[3] = make_float4(hi[0][0], hi[1][0], hi[2][0], hi[3][0]);
        o[4] = make_float4(hi[0][1], hi[1][1], hi[2][1], hi[3][1]);
        o[5] = make_float4(hi[0][2], hi[1][2], hi[2][2], hi[3][2]);
        *reinterpret_cast<int4 *>(o + 6) = make_int4(ref[0], ref[1], ref[2], ref[3]);
        o[7] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    }
}

template <bool COUNT>
__device__ __forceinline__ void bvh_closest(const SceneDev &S, f3 o, f3 d, Hit &h, uint32_t &nTri, uint32_t &nBox) {
    BvhWalk w;
    int2 stackMem[PTB_STACK];
    const LocalStack stack{stackMem};
    w.begin(S, o, d, h);
    while (w.round<COUNT>(S, stack, nTri, nBox)) {
    }
    h = w.h;
}

// ---- brute force: mode ALL (scene.cpp:77-86, object.cpp:19-34) ------------------------------
template <bool COUNT>
__device__ __forceinline__ void brute_closest(const SceneDev &S, f3 o, f3 d, Hit &h, uint32_t &nTri) {
    for (uint32_t i = 0; i < S.n_tris; ++i) intersect_tri<COUNT>(S.tri, (int)i, o, d, h, nTri);
}

// ---- the reference's tree, the reference's order: findBVHIntesection (bvh.cpp:48-99) statement by statement.
// Every popped node tests its own box (1 AS test), an inner node then measures both children (2 AS tests) and
// pushes far-then-near; a scene-level leaf pushes the roots of its objects, an object-level leaf tests its
// triangles.  Slower than bvh_closest on purpose — it exists so that nAS / nTri equal ray.interSectionAS /
// ray.interSectionTests of the reference (tracer.cpp:127-132) for the same ray.
template <bool COUNT>
__device__ __forceinline__ void refbvh_closest(const SceneDev &S, f3 o, f3 d, Hit &h, uint32_t &nTri, uint32_t &nAS) {
    if (S.n_tris == 0 || S.ref_root < 0) return;
    const f3 inv = rcp3(d);
    int stack[PTB_STACK];
    int sp = 0;
    stack[sp++] = S.ref_root;
    while (sp > 0) {
        const int idx = stack[--sp];
        const float4 *n = S.ref_node + 3 * (size_t)idx;
        const float4 a = __ldg(n), b = __ldg(n + 1);
        const int4 c = __ldg(reinterpret_cast<const int4 *>(n + 2));
        if (COUNT) nAS++;
        float tn;
        if (!slab(mk3(a.x, a.y, a.z), mk3(b.x, b.y, b.z), o, inv, h.t, tn)) continue;
        if (!(c.x == -1 && c.y == -1)) {
            if (COUNT) nAS += 2;
            const float4 *l = S.ref_node + 3 * (size_t)c.x, *r = S.ref_node + 3 * (size_t)c.y;
            const float4 la = __ldg(l), lb = __ldg(l + 1), ra = __ldg(r), rb = __ldg(r + 1);
            float dL, dR;
            if (!slab(mk3(la.x, la.y, la.z), mk3(lb.x, lb.y, lb.z), o, inv, h.t, dL)) dL = PTB_INF;
            if (!slab(mk3(ra.x, ra.y, ra.z), mk3(rb.x, rb.y, rb.z), o, inv, h.t, dR)) dR = PTB_INF;
            if (dR >= dL) {
                if (dR != PTB_INF && sp < PTB_STACK) stack[sp++] = c.y;
                if (dL != PTB_INF && sp < PTB_STACK) stack[sp++] = c.x;
            } else {
                if (dL != PTB_INF && sp < PTB_STACK) stack[sp++] = c.x;
                if (dR != PTB_INF && sp < PTB_STACK) stack[sp++] = c.y;
            }
        } else {
            const int first = __float_as_int(a.w), last = __float_as_int(b.w);
            for (int i = first; i < last; ++i) {
                const uint32_t id = __ldg(S.ref_idx + i);
                if (c.z) intersect_tri<COUNT>(S.tri, (int)__ldg(S.ref_tri_pos + id), o, d, h, nTri);
                else if (sp < PTB_STACK) stack[sp++] = __ldg(S.ref_obj_root + id);
            }
        }
    }
}

// dominant axis: lightFieldGrid.cpp:79-95 (strict >, first wins)
__device__ __forceinline__ int dominant_axis(f3 d) {
    int axis = 0;
    float m = 0.0f;
    const float f0 = fabsf(d.x), f1 = fabsf(d.y), f2 = fabsf(d.z);
    if (f0 > m) { m = f0; axis = 0; }
    if (f1 > m) { m = f1; axis = 1; }
    if (f2 > m) { m = f2; axis = 2; }
    return axis;
}

// ---- Subgrid ("2Plane"/Hybrid) closest hit: stands in for intersectGrid (lightFieldGrid.cpp:71-147)
// Grid header (4 float4): (aabb.min, size | hasTris<<31) (aabb.max, lutBase) (min, -) (inv_delta, -)
// grid.max[axis] is aabb.max[axis] + 0.1f (adjustGridSize, lightFieldGrid.cpp:176-178).
// The reference keeps a LIFO of grid ids and pushes every child of the cell it lands in (lightFieldGrid.cpp:127-145: ascending
// for a negative dominant direction, descending otherwise, so that pops come back-to-front resp. front-to-back).  A cell's
// children are one contiguous slice of the index pool, so the same pop order falls out of a LIFO of SLICES consumed from one
// end: one 8-byte push per visited inner grid instead of one load + store per child (14 % of the issued instructions of the
// round-1 kernel at 3 active lanes), and a stack as deep as the grid hierarchy — at most 1 + ceil(BVH depth / 4) <= 17 slices,
// a bound the builders check (ptb_build_subgrid[_gpu]) and ptb_upload_subgrid validates — instead of the 160 ids of round 1's
// walk, which dropped pushes silently when full.
#define PTB_GDEPTH 20
template <bool COUNT>
__device__ __forceinline__ void subgrid_closest_slices(const SceneDev &S, f3 o, f3 d, Hit &h, uint32_t &nTri, uint32_t &nAS) {
    const f3 inv = rcp3(d);
    const int axis = dominant_axis(d);
    const int right = axis == 0 ? 1 : 0, up = axis == 2 ? 1 : 2; // lightFieldGrid.cpp:26-35
    const float oa = comp(o, axis), ia = comp(inv, axis), da = comp(d, axis);
    const float orr = comp(o, right), dr = comp(d, right), ou = comp(o, up), du = comp(d, up);
    const bool fromEnd = da < 0.0f; // ascending pushes pop from the end of the slice
    const bool fast = finite3(inv);  // FMNMX slab test: same boolean as the select form for a finite inv_dir (see slab_fast)
    uint2 stack[PTB_GDEPTH];
    int sp = 0;
    uint32_t g = (uint32_t)axis; // grids 0..2 are the roots per dominant axis
    for (;;) {
        const float4 *hd = S.grid_hdr + 4 * (size_t)g;
        const float4 h0 = __ldg(hd), h1 = __ldg(hd + 1);
        if (COUNT) nAS++;
        float tn;
        if (fast ? slab_fast(mk3(h0.x, h0.y, h0.z), mk3(h1.x, h1.y, h1.z), o, inv, h.t, tn)
                 : slab(mk3(h0.x, h0.y, h0.z), mk3(h1.x, h1.y, h1.z), o, inv, h.t, tn)) {
            if (COUNT) nAS++;
            const float4 h2 = __ldg(hd + 2), h3 = __ldg(hd + 3);
            const f3 gmin = mk3(h2.x, h2.y, h2.z), idl = mk3(h3.x, h3.y, h3.z);
            const float gmaxA = comp(mk3(h1.x, h1.y, h1.z), axis) + 0.1f;
            const float d1 = (comp(gmin, axis) - oa) * ia;
            const float d2 = (gmaxA - oa) * ia;
            // in = (origin + d1*dir - min) * inv_delta, out likewise (lightFieldGrid.cpp:116-117)
            const float inR = ((orr + d1 * dr) - comp(gmin, right)) * comp(idl, right);
            const float inU = ((ou + d1 * du) - comp(gmin, up)) * comp(idl, up);
            const float outR = ((orr + d2 * dr) - comp(gmin, right)) * comp(idl, right);
            const float outU = ((ou + d2 * du) - comp(gmin, up)) * comp(idl, up);
            const uint32_t sz = __float_as_uint(h0.w) & 0x7fffffffu;
            const bool hasTris = (__float_as_uint(h0.w) >> 31) != 0u;
            const float fs = (float)sz;
            const int iu = (int)(inR * fs), iv = (int)(inU * fs), is = (int)(outR * fs), it = (int)(outU * fs);
            // the reference indexes without a bounds check (A.12); outside the planes means no cell
            if ((unsigned)iu < sz && (unsigned)iv < sz && (unsigned)is < sz && (unsigned)it < sz) {
                const uint32_t cell = ((((uint32_t)iu * sz + (uint32_t)iv) * sz + (uint32_t)is) * sz + (uint32_t)it);
                const uint2 se = __ldg(&S.grid_lut[(size_t)__float_as_uint(h1.w) + cell]);
                if (hasTris) {
                    for (uint32_t i = se.x; i < se.y; ++i) intersect_tri<COUNT>(S.tri, (int)__ldg(&S.grid_idx[i]), o, d, h, nTri);
                } else if (se.x < se.y && sp < PTB_GDEPTH) {
                    stack[sp++] = se;
                }
            }
        }
        if (sp == 0) break;
        uint2 e = stack[sp - 1];
        g = __ldg(&S.grid_idx[fromEnd ? --e.y : e.x++]);
        if (e.x == e.y) sp--;
        else stack[sp - 1] = e;
    }
}

// The same walk as a resumable state machine for the path kernel (k_mega<TWO_PLANE>): "while-while" like BvhWalk::round.
//   phase 1  lanes without pending triangles visit grids (header + slab + LUT cell) until one lands in a triangle cell or
//            runs out of grids;
//   phase 2  lanes holding a triangle cell test up to PTB_GRID_TRIS_PER_ROUND of its triangles together.
// A cell's triangles are all tested BEFORE the next grid is popped (the shrunken tmax culls it), exactly the order of
// intersectGrid, so hits, AS / triangle test counts and heat maps are unchanged.  Why: ncu on the one-shot walk inside k_mega
// shows 6.4 of 32 lanes per issued instruction (profiles/r2_ncu_mega_lucy_subgrid.txt) — every lane runs its own mixture of
// header fetches and triangle loops (cells hold up to m = 50 triangles) to completion while the warp waits for the longest.
#ifndef PTB_GRID_TRIS_PER_ROUND
#define PTB_GRID_TRIS_PER_ROUND 4
#endif
struct GridWalk {
    f3 o, d, inv;
    Hit h;
    int ref;          // grid to visit next, or PTB_REF_DONE (same convention as BvhWalk: the kernel only compares with DONE)
    int sp;           // slices on the stack
    uint32_t ti, te;  // pending triangles of the current cell: grid_idx[ti .. te)
    int axis;
    bool fast;        // inv_dir finite: the FMNMX slab test gives the same boolean as the select form (see slab_fast)

    __device__ __forceinline__ void begin(const SceneDev &S, f3 o_, f3 d_, const Hit &h0) {
        o = o_;
        d = d_;
        inv = rcp3(d_);
        h = h0;
        sp = 0;
        ti = te = 0u;
        axis = dominant_axis(d_);
        fast = finite3(inv);
        ref = axis; // grids 0..2 are the roots per dominant axis
    }
    template <class ST>
    __device__ __forceinline__ void next(const SceneDev &S, const ST &stack) {
        if (sp == 0) {
            ref = PTB_REF_DONE;
            return;
        }
        int2 e = stack.get(sp - 1);
        const bool fromEnd = comp(d, axis) < 0.0f; // ascending pushes pop from the end of the slice (lightFieldGrid.cpp:127-145)
        ref = (int)__ldg(&S.grid_idx[fromEnd ? (uint32_t)--e.y : (uint32_t)e.x++]);
        if (e.x == e.y) sp--;
        else stack.put(sp - 1, e);
    }
    // (Measured and rejected, profiles/r2_ab5_seek_cell_spec.txt: splitting the visit into "seek the next grid whose box is hit" and
    // "cell lookup for the lanes that found one" — the cell lookup runs at 4.5 lanes because ~60 % of the popped grids fail their
    // box test — made the path 15 % SLOWER: the extra divergent loop and the two header float4 carried across it cost more.)
    template <bool COUNT, class ST>
    __device__ __forceinline__ void visit(const SceneDev &S, const ST &stack, uint32_t &nAS) {
        const float4 *hd = S.grid_hdr + 4 * (size_t)ref;
        const float4 h0 = __ldg(hd), h1 = __ldg(hd + 1);
        if (COUNT) nAS++;
        float tn;
        if (!(fast ? slab_fast(mk3(h0.x, h0.y, h0.z), mk3(h1.x, h1.y, h1.z), o, inv, h.t, tn)
                   : slab(mk3(h0.x, h0.y, h0.z), mk3(h1.x, h1.y, h1.z), o, inv, h.t, tn)))
            return;
        if (COUNT) nAS++;
        const float4 h2 = __ldg(hd + 2), h3 = __ldg(hd + 3);
        const int right = axis == 0 ? 1 : 0, up = axis == 2 ? 1 : 2; // lightFieldGrid.cpp:26-35
        const float oa = comp(o, axis), ia = comp(inv, axis);
        const float orr = comp(o, right), dr = comp(d, right), ou = comp(o, up), du = comp(d, up);
        const f3 gmin = mk3(h2.x, h2.y, h2.z), idl = mk3(h3.x, h3.y, h3.z);
        const float gmaxA = comp(mk3(h1.x, h1.y, h1.z), axis) + 0.1f;
        const float d1 = (comp(gmin, axis) - oa) * ia;
        const float d2 = (gmaxA - oa) * ia;
        const float inR = ((orr + d1 * dr) - comp(gmin, right)) * comp(idl, right);
        const float inU = ((ou + d1 * du) - comp(gmin, up)) * comp(idl, up);
        const float outR = ((orr + d2 * dr) - comp(gmin, right)) * comp(idl, right);
        const float outU = ((ou + d2 * du) - comp(gmin, up)) * comp(idl, up);
        const uint32_t sz = __float_as_uint(h0.w) & 0x7fffffffu;
        const bool hasTris = (__float_as_uint(h0.w) >> 31) != 0u;
        const float fs = (float)sz;
        const int iu = (int)(inR * fs), iv = (int)(inU * fs), is = (int)(outR * fs), it = (int)(outU * fs);
        if ((unsigned)iu >= sz || (unsigned)iv >= sz || (unsigned)is >= sz || (unsigned)it >= sz) return; // outside the planes: no cell (A.12)
        const uint32_t cell = ((((uint32_t)iu * sz + (uint32_t)iv) * sz + (uint32_t)is) * sz + (uint32_t)it);
        const uint2 se = __ldg(&S.grid_lut[(size_t)__float_as_uint(h1.w) + cell]);
        if (se.x >= se.y) return;
        if (hasTris) {
            ti = se.x;
            te = se.y;
        } else if (sp < PTB_GDEPTH) {
            stack.put(sp++, make_int2((int)se.x, (int)se.y));
        }
    }
    template <bool COUNT, bool EARLY = false, class ST>
    __device__ __forceinline__ bool round(const SceneDev &S, const ST &stack, uint32_t &nTri, uint32_t &nAS, int minActive = 0) {
        while (ti == te && ref != PTB_REF_DONE) {
            visit<COUNT>(S, stack, nAS);
            if (ti == te) next(S, stack);
            if (EARLY && __popc(__activemask()) < minActive) break;
        }
        if (ti != te) {
            const uint32_t stop = min(te, ti + (uint32_t)PTB_GRID_TRIS_PER_ROUND);
            for (; ti < stop; ++ti) intersect_tri<COUNT>(S.tri, (int)__ldg(&S.grid_idx[ti]), o, d, h, nTri);
            if (ti == te) next(S, stack);
        }
        return ref != PTB_REF_DONE || ti != te;
    }
};

// ---- Subbeam closest hit: stands in for intersectSubBeamGrid (SubBeams.cpp:73-170) ---------
template <bool COUNT>
__device__ __forceinline__ void subbeam_closest(const SceneDev &S, f3 o, f3 d, Hit &h, uint32_t &nTri, uint32_t &nAS) {
    const f3 inv = rcp3(d);
    const int axis = dominant_axis(d);
    const int right = axis == 0 ? 1 : 0, up = axis == 2 ? 1 : 2;
    const f3 gmin = mk3(S.beam_gmin[axis][0], S.beam_gmin[axis][1], S.beam_gmin[axis][2]);
    const f3 gmax = mk3(S.beam_gmax[axis][0], S.beam_gmax[axis][1], S.beam_gmax[axis][2]);
    if (COUNT) nAS++;
    const float oa = comp(o, axis), ia = comp(inv, axis);
    const float d1 = (comp(gmin, axis) - oa) * ia;
    const float d2 = (comp(gmax, axis) - oa) * ia;
    if (smax(d1, d2) < 0.0f) return;
    if (COUNT) nAS++;
    float tn;
    if (!slab(mk3(S.beam_amin[0], S.beam_amin[1], S.beam_amin[2]), mk3(S.beam_amax[0], S.beam_amax[1], S.beam_amax[2]), o, inv, h.t, tn)) return;
    const float orr = comp(o, right), dr = comp(d, right), ou = comp(o, up), du = comp(d, up);
    const float ext_r = comp(gmax, right) - comp(gmin, right), ext_u = comp(gmax, up) - comp(gmin, up);
    float uvst[4];
    uvst[0] = ((orr + d1 * dr) - comp(gmin, right)) / ext_r;
    uvst[1] = ((ou + d1 * du) - comp(gmin, up)) / ext_u;
    uvst[2] = ((orr + d2 * dr) - comp(gmin, right)) / ext_r;
    uvst[3] = ((ou + d2 * du) - comp(gmin, up)) / ext_u;
    const uint32_t gs = S.beam_grid_size;
    const float fs = (float)gs;
    const int iu = (int)(uvst[0] * fs), iv = (int)(uvst[1] * fs), is = (int)(uvst[2] * fs), it = (int)(uvst[3] * fs);
    if ((unsigned)iu >= gs || (unsigned)iv >= gs || (unsigned)is >= gs || (unsigned)it >= gs) return;
    uint32_t b = __ldg(&S.beam_root[axis][(((uint32_t)iu * gs + (uint32_t)iv) * gs + (uint32_t)is) * gs + (uint32_t)it]);
    const float invg = 1.0f / fs;
    for (;;) {
        if (b >= S.n_beams) return;
        const float4 mn = __ldg(&S.beam[3 * (size_t)b]);
        const float4 mx = __ldg(&S.beam[3 * (size_t)b + 1]);
        const uint4 se = __ldg(reinterpret_cast<const uint4 *>(&S.beam[3 * (size_t)b + 2]));
        if (se.x >= se.y) return;
        if (se.z) {
            for (uint32_t i = se.x; i < se.y; ++i) intersect_tri<COUNT>(S.tri, (int)__ldg(&S.beam_idx[i]), o, d, h, nTri);
            return;
        }
        if (COUNT) nAS++;
        // scaled = (ray - min/gridSize) / ((max-min)/gridSize); Vector4/float is v*(1.0f/s) (vector.cpp)
        const float s0 = (uvst[0] - mn.x * invg) / ((mx.x - mn.x) * invg);
        const float s1 = (uvst[1] - mn.y * invg) / ((mx.y - mn.y) * invg);
        const float s2 = (uvst[2] - mn.z * invg) / ((mx.z - mn.z) * invg);
        const float s3 = (uvst[3] - mn.w * invg) / ((mx.w - mn.w) * invg);
        if (s0 < 0 || s0 > 1 || s1 < 0 || s1 > 1 || s2 < 0 || s2 > 1 || s3 < 0 || s3 > 1) return;
        // getBeamIdx (SubBeams.cpp:28-34); a coordinate of exactly 1.0 would index past the 16
        // children there — clamped to the last child here (documented deviation, measure zero)
        const int c0 = min((int)floorf(s0 * 2.0f), 1), c1 = min((int)floorf(s1 * 2.0f), 1);
        const int c2 = min((int)floorf(s2 * 2.0f), 1), c3 = min((int)floorf(s3 * 2.0f), 1);
        b = se.x + (uint32_t)(c0 * 8 + c1 * 4 + c2 * 2 + c3);
    }
}

template <bool COUNT>
__device__ __forceinline__ void bvh4_closest(const SceneDev &S, f3 o, f3 d, Hit &h, uint32_t &nTri, uint32_t &nBox) {
    BvhWalk4 w;
    int2 stackMem[PTB_STACK];
    const LocalStack stack{stackMem};
    w.begin(S, o, d, h);
    while (w.round<COUNT>(S, stack, nTri, nBox)) {
    }
    h = w.h;
}

// MODE: ptb_mode 0..3; 4 = mode BVH over the reference's own tree; 5 = mode BVH over the 4-wide collapse of this library's tree
template <int MODE, bool COUNT>
__device__ __forceinline__ void closest_hit(const SceneDev &S, f3 o, f3 d, Hit &h, uint32_t &nTri, uint32_t &nAS) {
    if (MODE == 5) bvh4_closest<COUNT>(S, o, d, h, nTri, nAS);
    else if (MODE == 1) bvh_closest<COUNT>(S, o, d, h, nTri, nAS);
    else if (MODE == 2) subgrid_closest_slices<COUNT>(S, o, d, h, nTri, nAS);
    else if (MODE == 3) subbeam_closest<COUNT>(S, o, d, h, nTri, nAS);
    else if (MODE == 4) refbvh_closest<COUNT>(S, o, d, h, nTri, nAS); // BVH mode over the reference's own tree
    else brute_closest<COUNT>(S, o, d, h, nTri);
}

} // namespace ptb
