// ptb_grid_build.cpp — host builders for the Subgrid ("2Plane"/Hybrid) hierarchy and the Subbeams,
// written straight into the flattened device layouts (see ptb_grid_build.h).
//
// What is kept from the reference (it defines which rays find which triangles):
//   * the two enlarged planes per grid (adjustGridSize, lightFieldGrid.cpp:165-187)
//   * a channel = sheared cuboid between cell (u,v) on the front plane and (s,t) on the back
//     plane; a triangle / child box belongs to it iff no separating axis exists among the axes
//     the reference tests (sat.cpp:8-84 cuboid-vs-AABB, :177-257 triangle-vs-channel)
//   * grid hierarchy = 4 BVH levels per grid level, leaf grids of <= m triangles
//     (lightFieldGrid.cpp:216-294), children ordered by aabb.min[axis] (:395-400)
// What is new: one BVH over all triangles instead of scene+object trees, channel-side projections
// computed once per cell instead of once per candidate, grids of a level built in parallel.
#include "ptb_grid_build.h"
#include "ptb_grid_gpu.h"
#include "ptb_sat.h"

#include <algorithm>
#include <atomic>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <thread>

namespace ptb {
namespace {

template <typename F>
void parallel_for(size_t n, F f) {
    const unsigned T = std::max(1u, std::min<unsigned>(std::thread::hardware_concurrency(), (unsigned)n));
    if (T <= 1) {
        for (size_t i = 0; i < n; ++i) f(i);
        return;
    }
    std::atomic<size_t> next{0};
    std::vector<std::thread> th;
    for (unsigned k = 0; k < T; ++k)
        th.emplace_back([&]() {
            for (;;) {
                size_t i = next.fetch_add(1);
                if (i >= n) break;
                f(i);
            }
        });
    for (auto &t : th) t.join();
}

struct PendingGrid {
    int axis;
    int32_t ref;     // BVH reference this grid was created from
    V3 lo, hi;       // its box
    uint32_t first, count; // leaf-order triangle range
    // filled while building
    bool hasTris = false;
    uint32_t size = 0;
    V3 gmin, gmax;
    std::vector<uint32_t> cand; // child grid ids or triangle indices
    std::vector<GridCell> lut;  // local ranges
    std::vector<uint32_t> idx;
};

} // namespace

void build_subgrid(const float *tri_leaf, uint64_t n, const BvhResult &bvh, uint32_t gs, uint32_t os, uint32_t m, SubgridResult &out) {
    out = SubgridResult();
    std::vector<PendingGrid> G;
    V3 slo{bvh.lo[0], bvh.lo[1], bvh.lo[2]}, shi{bvh.hi[0], bvh.hi[1], bvh.hi[2]};
    if (n == 0) {
        slo = {0, 0, 0};
        shi = {0, 0, 0};
    }
    for (int a = 0; a < 3; ++a) {
        PendingGrid g;
        g.axis = a;
        g.ref = bvh.root;
        g.lo = slo;
        g.hi = shi;
        g.first = 0;
        g.count = (uint32_t)n;
        G.push_back(g);
    }
    auto child_info = [&](int32_t node, int which, int32_t &ref, V3 &lo, V3 &hi, uint32_t &first, uint32_t &count) {
        const BvhNode64 &nd = bvh.nodes[node];
        ref = which ? nd.c1 : nd.c0;
        const float *l = which ? nd.lo1 : nd.lo0, *h = which ? nd.hi1 : nd.hi0;
        lo = {l[0], l[1], l[2]};
        hi = {h[0], h[1], h[2]};
        first = bvh.ranges[2 * (size_t)node + which].first;
        count = bvh.ranges[2 * (size_t)node + which].count;
    };
    // breadth-first over grid levels; a level's grids are independent
    size_t levelBegin = 0;
    while (levelBegin < G.size()) {
        const size_t levelEnd = G.size();
        // 1. topology (serial, cheap): decide inner/leaf and create the children
        for (size_t gi = levelBegin; gi < levelEnd; ++gi) {
            const bool inner = G[gi].ref >= 0 && G[gi].count > m && n > 0;
            if (inner) {
                struct Sub {
                    int32_t ref;
                    V3 lo, hi;
                    uint32_t first, count;
                };
                std::vector<Sub> cur{{G[gi].ref, G[gi].lo, G[gi].hi, G[gi].first, G[gi].count}}, nxt;
                for (int lvl = 0; lvl < 4; ++lvl) { // 4 BVH levels per grid level (lightFieldGrid.cpp:221-251)
                    nxt.clear();
                    for (auto &s : cur) {
                        if (s.ref < 0) {
                            nxt.push_back(s);
                            continue;
                        }
                        for (int w = 0; w < 2; ++w) {
                            Sub c;
                            child_info(s.ref, w, c.ref, c.lo, c.hi, c.first, c.count);
                            nxt.push_back(c);
                        }
                    }
                    cur.swap(nxt);
                }
                G[gi].hasTris = false;
                G[gi].size = gs;
                for (auto &s : cur) {
                    PendingGrid c;
                    c.axis = G[gi].axis;
                    c.ref = s.ref;
                    c.lo = s.lo;
                    c.hi = s.hi;
                    c.first = s.first;
                    c.count = s.count;
                    G[gi].cand.push_back((uint32_t)G.size());
                    G.push_back(c);
                }
            } else {
                G[gi].hasTris = true;
                G[gi].size = os;
                G[gi].cand.resize(G[gi].count);
                for (uint32_t k = 0; k < G[gi].count; ++k) G[gi].cand[k] = G[gi].first + k;
            }
        }
        // 2. channel population (parallel, the expensive part)
        parallel_for(levelEnd - levelBegin, [&](size_t k) {
            PendingGrid &g = G[levelBegin + k];
            g.gmin = g.lo;
            g.gmax = g.hi;
            enlarge(g.gmin, g.gmax, g.axis);
            const uint32_t sz = g.size;
            const float step = 1.0f / (float)sz;
            g.lut.resize((size_t)sz * sz * sz * sz);
            struct Ord {
                uint32_t id;
                float d;
            };
            std::vector<Ord> ord;
            Channel ch;
            for (uint32_t u = 0; u < sz; ++u)
                for (uint32_t v = 0; v < sz; ++v)
                    for (uint32_t s = 0; s < sz; ++s)
                        for (uint32_t t = 0; t < sz; ++t) {
                            const float fu = (float)u / (float)sz, fv = (float)v / (float)sz, fs = (float)s / (float)sz, ft = (float)t / (float)sz;
                            make_channel(g.gmin, g.gmax, g.axis, fu, fv, fu + step, fv + step, fs, ft, fs + step, ft + step, ch);
                            const uint32_t start = (uint32_t)g.idx.size();
                            if (g.hasTris) {
                                for (uint32_t ti : g.cand)
                                    if (tri_in_channel(tri_leaf + 9 * (size_t)ti, ch)) g.idx.push_back(ti);
                            } else {
                                ord.clear();
                                for (uint32_t ci : g.cand)
                                    if (box_in_channel(G[ci].lo, G[ci].hi, ch)) ord.push_back({ci, G[ci].lo[g.axis]});
                                std::stable_sort(ord.begin(), ord.end(), [](const Ord &a, const Ord &b) { return a.d < b.d; });
                                for (auto &o : ord) g.idx.push_back(o.id);
                            }
                            g.lut[(((size_t)u * sz + v) * sz + s) * sz + t] = {start, (uint32_t)g.idx.size()};
                        }
        });
        levelBegin = levelEnd;
    }
    // 3. flatten into the pools
    out.hdr.resize(G.size());
    size_t lutTotal = 0, idxTotal = 0;
    for (auto &g : G) {
        lutTotal += g.lut.size();
        idxTotal += g.idx.size();
    }
    out.lut.resize(lutTotal);
    out.idx.resize(idxTotal);
    size_t lutAt = 0, idxAt = 0;
    for (size_t gi = 0; gi < G.size(); ++gi) {
        PendingGrid &g = G[gi];
        GridHdr &h = out.hdr[gi];
        for (int k = 0; k < 3; ++k) {
            h.amin[k] = g.lo[k];
            h.amax[k] = g.hi[k];
            h.gmin[k] = g.gmin[k];
            h.inv_delta[k] = 1.0f / (g.gmax[k] - g.gmin[k]); // lightFieldGrid.cpp:302-304
        }
        h.size_flags = g.size | (g.hasTris ? 0x80000000u : 0u);
        h.lut_base = (uint32_t)lutAt;
        h.pad0 = h.pad1 = 0;
        for (size_t c = 0; c < g.lut.size(); ++c) out.lut[lutAt + c] = {g.lut[c].start + (uint32_t)idxAt, g.lut[c].end + (uint32_t)idxAt};
        if (!g.idx.empty()) memcpy(&out.idx[idxAt], g.idx.data(), g.idx.size() * sizeof(uint32_t));
        lutAt += g.lut.size();
        idxAt += g.idx.size();
        std::vector<GridCell>().swap(g.lut);
        std::vector<uint32_t>().swap(g.idx);
    }
}

// The same hierarchy walk as step 1 of build_subgrid, written into flat arrays for the device builder
// (ptb_grid_gpu.cu): same grid numbering, same candidate order after the reference's sort by aabb.min[axis].
void subgrid_topology(uint64_t n, const BvhResult &bvh, uint32_t gs, uint32_t os, uint32_t m, SubgridTopo &out) {
    out = SubgridTopo();
    struct Pend {
        int axis;
        int32_t ref;
        V3 lo, hi;
        uint32_t first, count;
    };
    std::vector<Pend> G;
    V3 slo{bvh.lo[0], bvh.lo[1], bvh.lo[2]}, shi{bvh.hi[0], bvh.hi[1], bvh.hi[2]};
    if (n == 0) {
        slo = {0, 0, 0};
        shi = {0, 0, 0};
    }
    for (int a = 0; a < 3; ++a) G.push_back(Pend{a, bvh.root, slo, shi, 0u, (uint32_t)n});
    out.cell_base.push_back(0);
    for (size_t gi = 0; gi < G.size(); ++gi) { // breadth-first: children are appended in the order build_subgrid creates them
        const Pend g = G[gi];
        GridTopo t{};
        V3 gmin = g.lo, gmax = g.hi;
        enlarge(gmin, gmax, g.axis);
        for (int k = 0; k < 3; ++k) {
            t.lo[k] = g.lo[k];
            t.hi[k] = g.hi[k];
            t.gmin[k] = gmin[k];
            t.gmax[k] = gmax[k];
        }
        const bool inner = g.ref >= 0 && g.count > m && n > 0;
        uint32_t size;
        if (inner) {
            std::vector<Pend> cur{g}, nxt;
            for (int lvl = 0; lvl < 4; ++lvl) {
                nxt.clear();
                for (auto &s : cur) {
                    if (s.ref < 0) {
                        nxt.push_back(s);
                        continue;
                    }
                    const BvhNode64 &nd = bvh.nodes[s.ref];
                    for (int w = 0; w < 2; ++w) {
                        const float *l = w ? nd.lo1 : nd.lo0, *h = w ? nd.hi1 : nd.hi0;
                        const BvhRange &r = bvh.ranges[2 * (size_t)s.ref + w];
                        nxt.push_back(Pend{g.axis, w ? nd.c1 : nd.c0, V3{l[0], l[1], l[2]}, V3{h[0], h[1], h[2]}, r.first, r.count});
                    }
                }
                cur.swap(nxt);
            }
            size = gs;
            t.cand_base = (uint32_t)out.cand.size();
            t.cand_count = (uint32_t)cur.size();
            struct Ord {
                uint32_t id;
                float d;
            };
            std::vector<Ord> ord;
            for (auto &s : cur) {
                ord.push_back({(uint32_t)G.size(), s.lo[g.axis]});
                G.push_back(s);
            }
            std::stable_sort(ord.begin(), ord.end(), [](const Ord &a, const Ord &b) { return a.d < b.d; });
            for (auto &o : ord) out.cand.push_back(o.id);
        } else {
            size = os;
            t.cand_base = g.first;
            t.cand_count = g.count;
        }
        t.axis_size_flags = (uint32_t)g.axis | (size << 8) | (inner ? 0u : 0x80000000u);
        out.max_cand = std::max(out.max_cand, t.cand_count);
        out.grids.push_back(t);
        out.cell_base.push_back(out.cell_base.back() + (uint64_t)size * size * size * size);
    }
}

GridHdr subgrid_header(const GridTopo &g, uint64_t lut_base) {
    GridHdr h;
    for (int k = 0; k < 3; ++k) {
        h.amin[k] = g.lo[k];
        h.amax[k] = g.hi[k];
        h.gmin[k] = g.gmin[k];
        h.inv_delta[k] = 1.0f / (g.gmax[k] - g.gmin[k]); // lightFieldGrid.cpp:302-304
    }
    h.size_flags = ((g.axis_size_flags >> 8) & 0xffffu) | (g.axis_size_flags & 0x80000000u);
    h.lut_base = (uint32_t)lut_base;
    h.pad0 = h.pad1 = 0;
    return h;
}

// =============================================================================================
// Subbeams (lightFieldGridSubBeams.cpp:185-374)
// =============================================================================================
namespace {

struct LocalBeam {
    float mn[4], mx[4];
    uint32_t start = 0, end = 0; // children: local beam indices; leaf: local idx range
    bool hasTris = false;
};

struct BeamTree { // the subtree under one root cell
    std::vector<LocalBeam> beams;
    std::vector<uint32_t> idx;
};

void build_beam_tree(const float *tri_leaf, V3 gmin, V3 gmax, int axis, float gsf, uint32_t maxTris, int maxDepth,
                     const std::vector<uint32_t> &rootCand, const float mn0[4], BeamTree &T) {
    struct Work {
        uint32_t beam;
        int depth;
        std::vector<uint32_t> cand;
    };
    T.beams.clear();
    T.idx.clear();
    LocalBeam root;
    for (int k = 0; k < 4; ++k) {
        root.mn[k] = mn0[k];
        root.mx[k] = mn0[k] + 1.0f;
    }
    T.beams.push_back(root);
    std::vector<Work> stack;
    stack.push_back({0u, 0, rootCand});
    const float inv = 1.0f / gsf; // Vector4 / float is v * (1.0f/s) (src/types/vector.cpp)
    Channel ch;
    while (!stack.empty()) {
        Work w = std::move(stack.back());
        stack.pop_back();
        LocalBeam b = T.beams[w.beam];
        float mn[4], mx[4];
        for (int k = 0; k < 4; ++k) {
            mn[k] = b.mn[k] * inv;
            mx[k] = b.mx[k] * inv;
        }
        make_channel(gmin, gmax, axis, mn[0], mn[1], mx[0], mx[1], mn[2], mn[3], mx[2], mx[3], ch);
        std::vector<uint32_t> in;
        for (uint32_t ti : w.cand)
            if (tri_in_channel(tri_leaf + 9 * (size_t)ti, ch)) in.push_back(ti);
        if (in.size() < maxTris || (maxDepth > 0 && w.depth >= maxDepth)) { // SubBeams.cpp:286
            T.beams[w.beam].hasTris = true;
            T.beams[w.beam].start = (uint32_t)T.idx.size();
            T.idx.insert(T.idx.end(), in.begin(), in.end());
            T.beams[w.beam].end = (uint32_t)T.idx.size();
        } else { // 2^4 children, child index = ((i*2 + j)*2 + k)*2 + h (getBeamIdx, SubBeams.cpp:28-34)
            const uint32_t first = (uint32_t)T.beams.size();
            float delta[4];
            for (int k = 0; k < 4; ++k) delta[k] = b.mx[k] - b.mn[k];
            for (int c = 0; c < 16; ++c) {
                const int ijkh[4] = {(c >> 3) & 1, (c >> 2) & 1, (c >> 1) & 1, c & 1};
                LocalBeam nb;
                for (int k = 0; k < 4; ++k) {
                    nb.mn[k] = b.mn[k] + (float)ijkh[k] * 0.5f * delta[k];
                    nb.mx[k] = nb.mn[k] + delta[k] * 0.5f;
                }
                T.beams.push_back(nb);
            }
            T.beams[w.beam].hasTris = false;
            T.beams[w.beam].start = first;
            T.beams[w.beam].end = first + 16;
            for (int c = 0; c < 16; ++c) stack.push_back({first + (uint32_t)c, w.depth + 1, in});
        }
    }
}

} // namespace

void build_subbeams(const float *tri_leaf, uint64_t n, const float scene_lo[3], const float scene_hi[3], uint32_t gs, uint32_t m,
                    int max_depth, SubbeamResult &out) {
    out = SubbeamResult();
    out.grid_size = gs;
    const size_t cells = (size_t)gs * gs * gs * gs;
    V3 slo{scene_lo[0], scene_lo[1], scene_lo[2]}, shi{scene_hi[0], scene_hi[1], scene_hi[2]};
    if (n == 0) {
        slo = {0, 0, 0};
        shi = {0, 0, 0};
    }
    for (int k = 0; k < 3; ++k) {
        out.amin[k] = slo[k];
        out.amax[k] = shi[k];
    }
    // triangle boxes for the per-(u,v) prefilter
    std::vector<float> tb(6 * n);
    for (uint64_t i = 0; i < n; ++i) {
        const float *p = tri_leaf + 9 * i;
        for (int k = 0; k < 3; ++k) {
            tb[6 * i + k] = std::min(p[k], std::min(p[3 + k], p[6 + k]));
            tb[6 * i + 3 + k] = std::max(p[k], std::max(p[3 + k], p[6 + k]));
        }
    }
    out.beams.resize(3 * cells);
    for (int axis = 0; axis < 3; ++axis) {
        V3 gmin = slo, gmax = shi;
        enlarge(gmin, gmax, axis);
        for (int k = 0; k < 3; ++k) {
            out.gmin[axis][k] = gmin[k];
            out.gmax[axis][k] = gmax[k];
        }
        out.root[axis].resize(cells);
        int right, up;
        axes_of(axis, right, up);
        std::vector<BeamTree> trees(cells);
        const float gsf = (float)gs;
        // one task per front cell (u,v): prefilter once, then all (s,t)
        parallel_for((size_t)gs * gs, [&](size_t uv) {
            const uint32_t u = (uint32_t)(uv / gs), v = (uint32_t)(uv % gs);
            // conservative box of every channel leaving front cell (u,v): front cell + whole back plane
            const float dR = gmax[right] - gmin[right], dU = gmax[up] - gmin[up];
            const float pad = 1e-3f * (dR + dU);
            float lo[3], hi[3];
            lo[axis] = gmin[axis] - pad;
            hi[axis] = gmax[axis] + pad;
            lo[right] = gmin[right] - pad;
            hi[right] = gmax[right] + pad;
            lo[up] = gmin[up] - pad;
            hi[up] = gmax[up] + pad;
            (void)u;
            (void)v; // the back plane spans the whole lateral extent, so the union is the enlarged box
            std::vector<uint32_t> cand;
            for (uint64_t i = 0; i < n; ++i) {
                bool sep = false;
                for (int k = 0; k < 3; ++k) sep |= tb[6 * i + 3 + k] < lo[k] || tb[6 * i + k] > hi[k];
                if (!sep) cand.push_back((uint32_t)i);
            }
            // second-level prefilter per (u,v,s): channels to back column s; exact channel test follows
            for (uint32_t s = 0; s < gs; ++s) {
                const float f0 = gmin[right] + dR * ((float)u / gsf), f1 = gmin[right] + dR * ((float)(u + 1) / gsf);
                const float b0 = gmin[right] + dR * ((float)s / gsf), b1 = gmin[right] + dR * ((float)(s + 1) / gsf);
                const float rlo = std::min(f0, b0) - pad, rhi = std::max(f1, b1) + pad;
                std::vector<uint32_t> cs;
                for (uint32_t ti : cand)
                    if (!(tb[6 * (size_t)ti + 3 + right] < rlo || tb[6 * (size_t)ti + right] > rhi)) cs.push_back(ti);
                for (uint32_t t = 0; t < gs; ++t) {
                    const float mn0[4] = {(float)u, (float)v, (float)s, (float)t};
                    const size_t cell = (((size_t)u * gs + v) * gs + s) * gs + t;
                    build_beam_tree(tri_leaf, gmin, gmax, axis, gsf, m, max_depth, cs, mn0, trees[cell]);
                }
            }
        });
        // merge: root beams live at [axis*cells + cell]; descendants are appended
        for (size_t cell = 0; cell < cells; ++cell) {
            BeamTree &T = trees[cell];
            const uint32_t rootId = (uint32_t)(axis * cells + cell);
            const uint32_t descBase = (uint32_t)out.beams.size();
            const uint32_t idxBase = (uint32_t)out.idx.size();
            out.root[axis][cell] = rootId;
            for (size_t li = 0; li < T.beams.size(); ++li) {
                const LocalBeam &lb = T.beams[li];
                BeamRec r;
                memcpy(r.mn, lb.mn, 16);
                memcpy(r.mx, lb.mx, 16);
                r.has_tris = lb.hasTris ? 1u : 0u;
                r.pad = 0;
                if (lb.hasTris) {
                    r.start = lb.start + idxBase;
                    r.end = lb.end + idxBase;
                } else {
                    r.start = descBase + lb.start - 1; // local children are >= 1
                    r.end = descBase + lb.end - 1;
                }
                if (li == 0) out.beams[rootId] = r;
                else out.beams.push_back(r);
            }
            out.idx.insert(out.idx.end(), T.idx.begin(), T.idx.end());
            BeamTree().beams.swap(T.beams);
            std::vector<uint32_t>().swap(T.idx);
        }
    }
}

} // namespace ptb
