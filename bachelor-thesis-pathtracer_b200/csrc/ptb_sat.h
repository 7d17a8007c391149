// ptb_sat.h — the channel geometry and separating-axis tests of the light-field structures, shared by
// the host builders (ptb_grid_build.cpp) and the device builder (ptb_grid_gpu.cu) so that both evaluate
// the SAME fp32 expressions in the same order: compiled without contraction on either side
// (-ffp-contract=off / -fmad=false, IEEE sqrt and division) they fill the same channels bit for bit.
//
// Kept from the reference (it defines which rays find which triangles):
//   * the two enlarged planes per grid (adjustGridSize, lightFieldGrid.cpp:165-187)
//   * a channel = sheared cuboid between cell (u,v) on the front plane and (s,t) on the back plane
//     (lightFieldGrid.cpp:509-569); a triangle / child box belongs to it iff no separating axis exists
//     among the axes the reference tests (sat.cpp:8-84 cuboid-vs-AABB, :177-257 triangle-vs-channel)
#pragma once
#include <cmath>

#ifdef __CUDACC__
#define PTB_HD __host__ __device__ __forceinline__
#define PTB_UNROLL _Pragma("unroll")
#else
#define PTB_HD inline
#define PTB_UNROLL
#endif

namespace ptb {

struct V3 {
    float x, y, z;
    PTB_HD float &operator[](int i) { return (&x)[i]; }
    PTB_HD float operator[](int i) const { return (&x)[i]; }
};
PTB_HD V3 sub(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
PTB_HD float dot(V3 a, V3 b) { return a.x * b.x + (a.y * b.y + a.z * b.z); }
PTB_HD V3 cross(V3 a, V3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
PTB_HD V3 normalized(V3 v) {
    float z = v.x * v.x + (v.y * v.y + v.z * v.z);
    if (z > 0.0f) {
        float l = sqrtf(z);
        return {v.x / l, v.y / l, v.z / l};
    }
    return v;
}
PTB_HD float min2(float a, float b) { return (b < a) ? b : a; } // std::min
PTB_HD float max2(float a, float b) { return (a < b) ? b : a; } // std::max

PTB_HD void axes_of(int axis, int &right, int &up) { // getGridAxes, lightFieldGrid.cpp:26-35
    right = axis == 0 ? 1 : 0;
    up = axis == 2 ? 1 : 2;
}

// adjustGridSize (lightFieldGrid.cpp:165-187 / lightFieldGridSubBeams.cpp:47-68)
PTB_HD void enlarge(V3 &gmin, V3 &gmax, int axis) {
    int right, up;
    axes_of(axis, right, up);
    const float offset = 0.1f;
    gmin[axis] -= offset;
    gmax[axis] += offset;
    const float deltaF = gmax[axis] - gmin[axis];
    gmin[up] -= offset + deltaF;
    gmin[right] -= offset + deltaF;
    gmax[up] += offset + deltaF;
    gmax[right] += offset + deltaF;
}

// A channel / beam: 8 corners, 3 edge directions, 3 face normals (lightFieldGrid.cpp:509-569) plus
// its own projection intervals on the axes that do not depend on the candidate.
struct Channel {
    V3 p[8], e[3], n[3];
    float cmin[3], cmax[3];   // coordinate axes
    float emin[3], emax[3];   // own edges
    float nmin[3], nmax[3];   // own normals
};

template <int N>
PTB_HD void interval(const V3 *pts, V3 ax, float &mn, float &mx) {
    mn = INFINITY;
    mx = -INFINITY;
    PTB_UNROLL
    for (int j = 0; j < N; ++j) {
        float pr = dot(pts[j], ax);
        if (pr < mn) mn = pr;
        if (pr > mx) mx = pr;
    }
}

// AXIS is a template parameter so that every component index is a constant (registers on the device)
template <int AXIS>
PTB_HD void make_channel_t(V3 gmin, V3 gmax, float u0, float v0, float u1, float v1, float s0, float t0, float s1, float t1, Channel &c) {
    constexpr int right = AXIS == 0 ? 1 : 0, up = AXIS == 2 ? 1 : 2;
    const float dR = gmax[right] - gmin[right], dU = gmax[up] - gmin[up];
#define PTB_SET(i, a, r, u)                 \
    c.p[i][AXIS] = (a);                     \
    c.p[i][right] = gmin[right] + dR * (r); \
    c.p[i][up] = gmin[up] + dU * (u);
    PTB_SET(0, gmin[AXIS], u0, v0)
    PTB_SET(1, gmin[AXIS], u1, v1)
    PTB_SET(2, gmax[AXIS], s0, t1)
    PTB_SET(3, gmax[AXIS], s1, t0)
    PTB_SET(4, gmin[AXIS], u1, v0)
    PTB_SET(5, gmin[AXIS], u0, v1)
    PTB_SET(6, gmax[AXIS], s1, t1)
    PTB_SET(7, gmax[AXIS], s0, t0)
#undef PTB_SET
    c.e[0] = normalized(sub(c.p[6], c.p[1]));
    c.e[1] = normalized(sub(c.p[2], c.p[6]));
    c.e[2] = normalized(sub(c.p[6], c.p[3]));
    c.n[0] = normalized(cross(c.e[0], c.e[1]));
    c.n[1] = normalized(cross(c.e[0], c.e[2]));
    c.n[2] = normalized(cross(c.e[2], c.e[1]));
    PTB_UNROLL
    for (int k = 0; k < 3; ++k) {
        c.cmin[k] = INFINITY;
        c.cmax[k] = -INFINITY;
        PTB_UNROLL
        for (int i = 0; i < 8; ++i) {
            if (c.p[i][k] < c.cmin[k]) c.cmin[k] = c.p[i][k];
            if (c.p[i][k] > c.cmax[k]) c.cmax[k] = c.p[i][k];
        }
        interval<8>(c.p, c.e[k], c.emin[k], c.emax[k]);
        interval<8>(c.p, c.n[k], c.nmin[k], c.nmax[k]);
    }
}

PTB_HD void make_channel(V3 gmin, V3 gmax, int axis, float u0, float v0, float u1, float v1, float s0, float t0, float s1, float t1, Channel &c) {
    if (axis == 0) make_channel_t<0>(gmin, gmax, u0, v0, u1, v1, s0, t0, s1, t1, c);
    else if (axis == 1) make_channel_t<1>(gmin, gmax, u0, v0, u1, v1, s0, t0, s1, t1, c);
    else make_channel_t<2>(gmin, gmax, u0, v0, u1, v1, s0, t0, s1, t1, c);
}

PTB_HD bool separated(float amin, float amax, float bmin, float bmax) { return amax < bmin || amin > bmax; }

// triangle vs channel: the 3 + 19 axes of triInChannel (sat.cpp:177-257)
PTB_HD bool tri_in_channel(const float *t9, const Channel &c) {
    const V3 tv[3] = {{t9[0], t9[1], t9[2]}, {t9[3], t9[4], t9[5]}, {t9[6], t9[7], t9[8]}};
    PTB_UNROLL
    for (int k = 0; k < 3; ++k) {
        float mn = min2(tv[0][k], min2(tv[1][k], tv[2][k])), mx = max2(tv[0][k], max2(tv[1][k], tv[2][k]));
        if (separated(mn, mx, c.cmin[k], c.cmax[k])) return false;
    }
    const V3 te[3] = {sub(tv[1], tv[0]), sub(tv[2], tv[1]), sub(tv[0], tv[2])};
    float a0, a1, b0, b1;
    PTB_UNROLL
    for (int i = 0; i < 3; ++i) {
        PTB_UNROLL
        for (int j = 0; j < 3; ++j) {
            V3 ax = cross(te[i], c.e[j]);
            interval<3>(tv, ax, a0, a1);
            interval<8>(c.p, ax, b0, b1);
            if (separated(a0, a1, b0, b1)) return false;
        }
    }
    PTB_UNROLL
    for (int i = 0; i < 3; ++i) {
        interval<3>(tv, te[i], a0, a1);
        interval<8>(c.p, te[i], b0, b1);
        if (separated(a0, a1, b0, b1)) return false;
    }
    {
        V3 ax = normalized(cross(te[0], te[1]));
        interval<3>(tv, ax, a0, a1);
        interval<8>(c.p, ax, b0, b1);
        if (separated(a0, a1, b0, b1)) return false;
    }
    PTB_UNROLL
    for (int i = 0; i < 3; ++i) {
        interval<3>(tv, c.e[i], a0, a1);
        if (separated(a0, a1, c.emin[i], c.emax[i])) return false;
    }
    PTB_UNROLL
    for (int i = 0; i < 3; ++i) {
        interval<3>(tv, c.n[i], a0, a1);
        if (separated(a0, a1, c.nmin[i], c.nmax[i])) return false;
    }
    return true;
}

// channel vs axis-aligned box: the 3 + 15 axes of cuboidInAABB (sat.cpp:8-84)
PTB_HD bool box_in_channel(V3 mn, V3 mx, const Channel &c) {
    PTB_UNROLL
    for (int k = 0; k < 3; ++k)
        if (separated(c.cmin[k], c.cmax[k], mn[k], mx[k])) return false;
    const V3 bv[8] = {{mn.x, mn.y, mn.z}, {mn.x, mn.y, mx.z}, {mn.x, mx.y, mn.z}, {mn.x, mx.y, mx.z},
                      {mx.x, mn.y, mn.z}, {mx.x, mn.y, mx.z}, {mx.x, mx.y, mn.z}, {mx.x, mx.y, mx.z}};
    const V3 ea[3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
    float a0, a1, b0, b1;
    PTB_UNROLL
    for (int i = 0; i < 3; ++i) {
        PTB_UNROLL
        for (int j = 0; j < 3; ++j) {
            V3 ax = cross(ea[i], c.e[j]);
            interval<8>(c.p, ax, a0, a1);
            interval<8>(bv, ax, b0, b1);
            if (separated(a0, a1, b0, b1)) return false;
        }
    }
    PTB_UNROLL
    for (int i = 0; i < 3; ++i) {
        interval<8>(bv, c.e[i], b0, b1);
        if (separated(c.emin[i], c.emax[i], b0, b1)) return false;
    }
    PTB_UNROLL
    for (int i = 0; i < 3; ++i) {
        interval<8>(bv, c.n[i], b0, b1);
        if (separated(c.nmin[i], c.nmax[i], b0, b1)) return false;
    }
    return true;
}

} // namespace ptb
