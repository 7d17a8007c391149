// ptb_beam_levels.h — the Subbeam structure built level by level (second half).
// Stands in for buildGridsWithSubBeams / the beam recursion
// (reference: src/accelerationStructures/lightFieldGridSubBeams.cpp:185-374).
//
// The reference (and the host builder, ptb_grid_build.cpp: build_beam_tree) recurse depth first: a beam tests its parent's
// accepted triangles against its own sheared cuboid, becomes a leaf when fewer than `maxTris` remain (or the depth guard
// fires) and otherwise splits into 2^4 children.  Every beam of one depth is independent of the others, so here a whole
// depth is one data-parallel step —
//   count   one worker per beam: SAT tests of its channel against the parent's list      (BeamCount<false>)
//           (depth 0, where every root tests every triangle: one worker per root and 1024 triangles, BeamRootChunks)
//   scan    exclusive sum = the beam's slice of this level's pool
//   fill    the accepted triangle ids, ascending like the host's lists                     (BeamCount<true>)
//   split   leaf / inner, inner beams ranked by a scan, 16 children each                   (BeamClassify, BeamChildren)
// — and the depth-first NUMBERING the traversal kernel and the host builder share (children numbered when their parent is
// popped from a LIFO stack, i.e. siblings 15..0, leaf slices in the same order) is recovered afterwards from subtree sizes:
//   up      subtree totals (inner beams, leaf entries) from the deepest level to the roots (BeamUp)
//   down    what was processed before each beam = parent's prefix + the later siblings' subtrees (BeamDown)
//   emit    48-byte beam records and leaf slices at their final places                      (BeamEmit)
// so the pools equal build_subbeams' byte for byte.
//
// The driver below is a template over an executor (allocate / run a functor over [0,n) / exclusive scan / read one value):
// ptb_beam_gpu.cu instantiates it with CUDA kernels and cub, tests/native/subbeam_levels_check.cpp with plain loops — the
// same functors and the same driver are checked against build_subbeams on the CPU before they ever run on a GPU.
#pragma once
#include <cstdint>
#include <vector>

#include "ptb_sat.h"

namespace ptb {

struct BeamLevelDev { // one depth of all beam trees, structure of arrays (executor memory)
    uint32_t n = 0;
    float *mn = nullptr, *mx = nullptr; // 4 floats per beam: uvst box in grid-cell units
    uint32_t *tree = nullptr;           // axis * cells + cell: which root this beam descends from
    uint32_t *parent = nullptr;         // index in the level above (depth > 0)
    uint32_t *cnt = nullptr, *start = nullptr, *pool = nullptr; // accepted triangles: pool[start .. start+cnt)
    uint32_t *inner = nullptr;          // 1: splits into 16 children at childBase .. childBase+15 of the next level
    uint32_t *childBase = nullptr;
    uint32_t *subInner = nullptr, *subIdx = nullptr;       // subtree totals: inner beams, leaf entries
    uint32_t *innerBefore = nullptr, *idxBefore = nullptr; // processed before this beam in the host's depth-first order
    uint64_t poolSize = 0;
};

struct BeamGeom { // per dominant axis: the enlarged scene box (adjustGridSize), as build_subbeams computes it
    float gmin[3][3], gmax[3][3];
    uint32_t gs = 0, cells = 0; // gridSize, gridSize^4
    float gsf = 0.0f, inv = 0.0f; // (float)gs, 1.0f / gsf
};

// ---- functors (one element each; PTB_HD so that host loops and CUDA kernels run the same code) ---------------------
struct BeamRootInit { // the 3 * gs^4 root beams: cell (u,v,s,t), one grid cell wide (build_subbeams: mn0 / mn0 + 1)
    BeamLevelDev L;
    uint32_t gs, cells;
    PTB_HD void operator()(uint32_t b) const {
        const uint32_t cell = b % cells;
        const uint32_t t = cell % gs, s = (cell / gs) % gs, v = (cell / (gs * gs)) % gs, u = cell / (gs * gs * gs);
        const float m[4] = {(float)u, (float)v, (float)s, (float)t};
        for (int k = 0; k < 4; ++k) {
            L.mn[4 * (size_t)b + k] = m[k];
            L.mx[4 * (size_t)b + k] = m[k] + 1.0f;
        }
        L.tree[b] = b;
    }
};

template <bool FILL>
struct BeamCount { // SAT tests of beam b's channel against its candidates (depth 0: every triangle; else the parent's list)
    BeamLevelDev L, P; // this level, the level above
    BeamGeom G;
    int depth;
    uint32_t nTris;
    const float *tri9;      // 9 floats per triangle
    const uint32_t *remap;  // leaf position -> row of tri9 (null: identity)
    PTB_HD void operator()(uint32_t b) const {
        const int axis = (int)(L.tree[b] / G.cells);
        float mn[4], mx[4];
        for (int k = 0; k < 4; ++k) {
            mn[k] = L.mn[4 * (size_t)b + k] * G.inv;
            mx[k] = L.mx[4 * (size_t)b + k] * G.inv;
        }
        Channel ch;
        make_channel(V3{G.gmin[axis][0], G.gmin[axis][1], G.gmin[axis][2]}, V3{G.gmax[axis][0], G.gmax[axis][1], G.gmax[axis][2]}, axis,
                     mn[0], mn[1], mx[0], mx[1], mn[2], mn[3], mx[2], mx[3], ch);
        uint32_t nc = nTris;
        const uint32_t *cand = nullptr;
        if (depth > 0) {
            const uint32_t p = L.parent[b];
            nc = P.cnt[p];
            cand = P.pool + P.start[p];
        }
        uint32_t *out = FILL ? L.pool + L.start[b] : nullptr;
        uint32_t c = 0;
        for (uint32_t k = 0; k < nc; ++k) {
            const uint32_t ti = cand ? cand[k] : k;
            const float *q = tri9 + 9 * (size_t)(remap ? remap[ti] : ti);
            float t9[9];
            PTB_UNROLL
            for (int j = 0; j < 9; ++j) t9[j] = q[j];
            if (tri_in_channel(t9, ch)) {
                if (FILL) out[c] = ti;
                c++;
            }
        }
        if (!FILL) L.cnt[b] = c;
    }
};

template <bool FILL>
struct BeamRootChunks { // depth 0 only: every root tests every triangle, so beam b's loop is cut into `nChunks` workers
    BeamLevelDev L;     // of `chunk` consecutive triangles each; slot b * nChunks + c keeps the ascending order under one scan.
    BeamGeom G;         // Worker i -> (b = i % L.n, c = i / L.n): neighbouring workers share their triangles (broadcast loads)
    uint32_t nTris, nChunks, chunk;
    const float *tri9;
    const uint32_t *remap;
    uint32_t *slotCnt;          // [L.n * nChunks + 1], the last one stays 0
    const uint32_t *slotStart;  // its exclusive scan (FILL)
    PTB_HD void operator()(uint32_t i) const {
        const uint32_t b = i % L.n, c = i / L.n;
        const int axis = (int)(L.tree[b] / G.cells);
        float mn[4], mx[4];
        for (int k = 0; k < 4; ++k) {
            mn[k] = L.mn[4 * (size_t)b + k] * G.inv;
            mx[k] = L.mx[4 * (size_t)b + k] * G.inv;
        }
        Channel ch;
        make_channel(V3{G.gmin[axis][0], G.gmin[axis][1], G.gmin[axis][2]}, V3{G.gmax[axis][0], G.gmax[axis][1], G.gmax[axis][2]}, axis,
                     mn[0], mn[1], mx[0], mx[1], mn[2], mn[3], mx[2], mx[3], ch);
        const size_t slot = (size_t)b * nChunks + c;
        const uint32_t k0 = c * chunk, k1 = (nTris - k0 < chunk) ? nTris : k0 + chunk;
        uint32_t *out = FILL ? L.pool + slotStart[slot] : nullptr;
        uint32_t n = 0;
        for (uint32_t ti = k0; ti < k1; ++ti) {
            const float *q = tri9 + 9 * (size_t)(remap ? remap[ti] : ti);
            float t9[9];
            PTB_UNROLL
            for (int j = 0; j < 9; ++j) t9[j] = q[j];
            if (tri_in_channel(t9, ch)) {
                if (FILL) out[n] = ti;
                n++;
            }
        }
        if (!FILL) slotCnt[slot] = n;
    }
};

struct BeamRootGather { // a root's list = its chunks' slots, contiguous after the scan
    BeamLevelDev L;
    uint32_t nChunks;
    const uint32_t *slotStart;
    PTB_HD void operator()(uint32_t b) const {
        const uint32_t s = slotStart[(size_t)b * nChunks];
        L.start[b] = s;
        L.cnt[b] = slotStart[(size_t)(b + 1) * nChunks] - s;
    }
};

// ---- depth >= 1, opt-in: a beam's walk over its parent's list cut into workers of `chunk` entries ------------------------
struct BeamSlotCount { // how many workers beam b needs
    BeamLevelDev L, P;
    uint32_t chunk;
    uint32_t *nSlots;
    PTB_HD void operator()(uint32_t b) const { nSlots[b] = (P.cnt[L.parent[b]] + chunk - 1) / chunk; }
};

template <bool FILL>
struct BeamSlots { // worker i -> (beam b with slotBase[b] <= i < slotBase[b + 1], chunk c = i - slotBase[b])
    BeamLevelDev L, P;
    BeamGeom G;
    uint32_t chunk;
    const float *tri9;
    const uint32_t *remap;
    const uint32_t *slotBase;   // [L.n + 1], exclusive scan of nSlots
    uint32_t *slotCnt;          // [slots + 1]
    const uint32_t *slotStart;  // its exclusive scan (FILL)
    PTB_HD void operator()(uint32_t i) const {
        uint32_t lo = 0, hi = L.n; // the last b with slotBase[b] <= i (beams without workers share their successor's base)
        while (hi - lo > 1) {
            const uint32_t mid = lo + (hi - lo) / 2;
            if (slotBase[mid] <= i) lo = mid;
            else hi = mid;
        }
        const uint32_t b = lo, c = i - slotBase[b];
        const int axis = (int)(L.tree[b] / G.cells);
        float mn[4], mx[4];
        for (int k = 0; k < 4; ++k) {
            mn[k] = L.mn[4 * (size_t)b + k] * G.inv;
            mx[k] = L.mx[4 * (size_t)b + k] * G.inv;
        }
        Channel ch;
        make_channel(V3{G.gmin[axis][0], G.gmin[axis][1], G.gmin[axis][2]}, V3{G.gmax[axis][0], G.gmax[axis][1], G.gmax[axis][2]}, axis,
                     mn[0], mn[1], mx[0], mx[1], mn[2], mn[3], mx[2], mx[3], ch);
        const uint32_t p = L.parent[b], total = P.cnt[p];
        const uint32_t *cand = P.pool + P.start[p];
        const uint32_t k0 = c * chunk, k1 = (total - k0 < chunk) ? total : k0 + chunk;
        uint32_t *out = FILL ? L.pool + slotStart[i] : nullptr;
        uint32_t n = 0;
        for (uint32_t k = k0; k < k1; ++k) {
            const uint32_t ti = cand[k];
            const float *q = tri9 + 9 * (size_t)(remap ? remap[ti] : ti);
            float t9[9];
            PTB_UNROLL
            for (int j = 0; j < 9; ++j) t9[j] = q[j];
            if (tri_in_channel(t9, ch)) {
                if (FILL) out[n] = ti;
                n++;
            }
        }
        if (!FILL) slotCnt[i] = n;
    }
};

struct BeamSlotGather {
    BeamLevelDev L;
    const uint32_t *slotBase, *slotStart;
    PTB_HD void operator()(uint32_t b) const {
        const uint32_t s = slotStart[slotBase[b]];
        L.start[b] = s;
        L.cnt[b] = slotStart[slotBase[b + 1]] - s;
    }
};

struct BeamClassify { // SubBeams.cpp:286: leaf iff fewer than maxTris triangles remain (or the depth guard fires)
    BeamLevelDev L;
    uint32_t maxTris;
    int depth, maxDepth;
    PTB_HD void operator()(uint32_t b) const { L.inner[b] = (L.cnt[b] < maxTris || (maxDepth > 0 && depth >= maxDepth)) ? 0u : 1u; }
};

struct BeamChildBase { // childBase = 16 * (rank among the inner beams of the level); innerList[rank] = b
    BeamLevelDev L;
    const uint32_t *rank;
    uint32_t *innerList;
    PTB_HD void operator()(uint32_t b) const {
        L.childBase[b] = 16u * rank[b];
        if (L.inner[b]) innerList[rank[b]] = b;
    }
};

struct BeamChildren { // child index = ((i*2 + j)*2 + k)*2 + h (getBeamIdx, SubBeams.cpp:28-34); halves of the parent's box
    BeamLevelDev C, P;
    const uint32_t *innerList;
    PTB_HD void operator()(uint32_t j) const {
        const uint32_t p = innerList[j >> 4];
        const int c = (int)(j & 15u);
        const int ijkh[4] = {(c >> 3) & 1, (c >> 2) & 1, (c >> 1) & 1, c & 1};
        for (int k = 0; k < 4; ++k) {
            const float pm = P.mn[4 * (size_t)p + k];
            const float delta = P.mx[4 * (size_t)p + k] - pm;
            const float m = pm + (float)ijkh[k] * 0.5f * delta;
            C.mn[4 * (size_t)j + k] = m;
            C.mx[4 * (size_t)j + k] = m + delta * 0.5f;
        }
        C.tree[j] = P.tree[p];
        C.parent[j] = p;
    }
};

struct BeamUp { // subtree totals; N = the level below (unused for leaves)
    BeamLevelDev L, N;
    PTB_HD void operator()(uint32_t b) const {
        if (!L.inner[b]) {
            L.subInner[b] = 0;
            L.subIdx[b] = L.cnt[b];
            return;
        }
        uint32_t si = 1, sx = 0;
        const uint32_t cb = L.childBase[b];
        for (uint32_t c = 0; c < 16; ++c) {
            si += N.subInner[cb + c];
            sx += N.subIdx[cb + c];
        }
        L.subInner[b] = si;
        L.subIdx[b] = sx;
    }
};

struct BeamDown { // the host pops siblings 15..0: before child c come the parent itself and the subtrees of c+1..15
    BeamLevelDev C, P;
    PTB_HD void operator()(uint32_t j) const {
        const uint32_t p = C.parent[j];
        const uint32_t c = j & 15u, base = j - c;
        uint32_t ib = P.innerBefore[p] + 1u, xb = P.idxBefore[p];
        for (uint32_t s = c + 1; s < 16; ++s) {
            ib += C.subInner[base + s];
            xb += C.subIdx[base + s];
        }
        C.innerBefore[j] = ib;
        C.idxBefore[j] = xb;
    }
};

struct BeamZeroBefore {
    BeamLevelDev L;
    PTB_HD void operator()(uint32_t b) const { L.innerBefore[b] = L.idxBefore[b] = 0; }
};

struct BeamRootTotals { // per tree: descendants (16 per inner beam) and leaf entries, for the scans that place the trees
    BeamLevelDev L;
    uint32_t *desc, *idx;
    PTB_HD void operator()(uint32_t b) const {
        desc[b] = 16u * L.subInner[b];
        idx[b] = L.subIdx[b];
    }
};

struct BeamEmit { // the 48-byte record (BeamRec: mn[4], mx[4], start, end, has_tris, pad) and the leaf slice at their final places
    BeamLevelDev L, P;
    int depth;
    uint32_t rootBeams;            // 3 * cells: roots live at [tree], descendants are appended tree after tree
    const uint32_t *descScan, *idxScan; // exclusive scans of the per-tree totals
    float *beams;                  // 12 x 4 bytes per record
    uint32_t *idx;
    PTB_HD void operator()(uint32_t b) const {
        const uint32_t tree = L.tree[b];
        const uint32_t descBase = rootBeams + descScan[tree], idxBase = idxScan[tree];
        uint32_t pos = tree; // local id 0
        if (depth > 0) {
            const uint32_t first = 1u + 16u * P.innerBefore[L.parent[b]]; // the parent's children were numbered when it was popped
            pos = descBase + (first + (b & 15u)) - 1u;
        }
        float *r = beams + 12 * (size_t)pos;
        uint32_t *ri = reinterpret_cast<uint32_t *>(r);
        for (int k = 0; k < 4; ++k) {
            r[k] = L.mn[4 * (size_t)b + k];
            r[4 + k] = L.mx[4 * (size_t)b + k];
        }
        if (L.inner[b]) {
            const uint32_t first = 1u + 16u * L.innerBefore[b];
            ri[8] = descBase + first - 1u;
            ri[9] = descBase + first - 1u + 16u;
            ri[10] = 0u;
        } else {
            const uint32_t s = idxBase + L.idxBefore[b];
            ri[8] = s;
            ri[9] = s + L.cnt[b];
            ri[10] = 1u;
            const uint32_t *src = L.pool + L.start[b];
            for (uint32_t k = 0; k < L.cnt[b]; ++k) idx[s + k] = src[k];
        }
        ri[11] = 0u;
    }
};

constexpr uint32_t kBeamRootChunk = 1024; // triangles per depth-0 worker

struct BeamLevelsResult {
    float *beams = nullptr;  // executor memory, 48 bytes per beam; owned by the caller (X::release)
    uint32_t *idx = nullptr;
    uint64_t nBeams = 0, nIdx = 0;
    int depth = 0;           // deepest level that holds beams
};

// Executor X: T *alloc<T>(count) (scratch, freed by X), T *alloc_out<T>(count) (results, owned by the caller),
// run(n, functor), scan(in, out, n) (exclusive, u32), sum64(in, n), fail() (a device error happened).
// Returns an empty string or what went wrong.
// deepChunk > 0 (opt-in): below the roots, too, a beam's list is walked by workers of `deepChunk` entries (BeamSlots).
template <class X>
const char *build_beam_levels(X &x, const BeamGeom &G, uint32_t nTris, const float *tri9, const uint32_t *remap, uint32_t maxTris,
                              int maxDepth, BeamLevelsResult &out, uint32_t deepChunk = 0) {
    const uint32_t roots = 3u * G.cells;
    std::vector<BeamLevelDev> lv;
    auto new_level = [&](uint32_t n) {
        BeamLevelDev L;
        L.n = n;
        L.mn = x.template alloc<float>(4 * (size_t)n);
        L.mx = x.template alloc<float>(4 * (size_t)n);
        L.tree = x.template alloc<uint32_t>(n);
        L.parent = x.template alloc<uint32_t>(n);
        L.cnt = x.template alloc<uint32_t>(n);
        L.start = x.template alloc<uint32_t>(n);
        L.inner = x.template alloc<uint32_t>(n);
        L.childBase = x.template alloc<uint32_t>(n);
        L.subInner = x.template alloc<uint32_t>(n);
        L.subIdx = x.template alloc<uint32_t>(n);
        L.innerBefore = x.template alloc<uint32_t>(n);
        L.idxBefore = x.template alloc<uint32_t>(n);
        return L;
    };
    lv.push_back(new_level(roots));
    x.run(roots, BeamRootInit{lv[0], G.gs, G.cells});
    uint64_t totalPool = 0;
    for (int d = 0;; ++d) {
        if (x.fail()) return "device error while building the beam levels";
        BeamLevelDev &L = lv[d];
        const BeamLevelDev P = d > 0 ? lv[d - 1] : BeamLevelDev();
        // depth 0 with more than one chunk of triangles: (root, chunk) workers, see BeamRootChunks
        uint32_t chunk = kBeamRootChunk;
        while (d == 0 && (uint64_t)roots * ((nTris + chunk - 1) / chunk) > (1ull << 28)) chunk *= 2;
        const uint32_t nChunks = d == 0 ? (nTris + chunk - 1) / chunk : 1;
        if (nChunks > 1) {
            const uint32_t slots = roots * nChunks;
            uint32_t *slotCnt = x.template alloc<uint32_t>((size_t)slots + 1), *slotStart = x.template alloc<uint32_t>((size_t)slots + 1);
            x.run(slots, BeamRootChunks<false>{L, G, nTris, nChunks, chunk, tri9, remap, slotCnt, slotStart});
            L.poolSize = x.sum64(slotCnt, slots);
            if (L.poolSize >= (1ull << 32)) return "Subbeam pools exceed 2^32 entries: lower beam_max_depth (ptb_set_build_limits) or the grid size";
            x.scan(slotCnt, slotStart, slots + 1);
            L.pool = x.template alloc<uint32_t>((size_t)L.poolSize);
            x.run(slots, BeamRootChunks<true>{L, G, nTris, nChunks, chunk, tri9, remap, slotCnt, slotStart});
            x.run(roots, BeamRootGather{L, nChunks, slotStart});
            totalPool += L.poolSize;
        } else if (d > 0 && deepChunk > 0) {
            uint32_t *nSlots = x.template alloc<uint32_t>((size_t)L.n + 1), *slotBase = x.template alloc<uint32_t>((size_t)L.n + 1);
            x.run(L.n, BeamSlotCount{L, P, deepChunk, nSlots});
            const uint64_t slots64 = x.sum64(nSlots, L.n);
            if (slots64 >= (1ull << 31)) return "Subbeam level needs more than 2^31 workers: raise the chunk size";
            const uint32_t slots = (uint32_t)slots64;
            x.scan(nSlots, slotBase, L.n + 1);
            uint32_t *slotCnt = x.template alloc<uint32_t>((size_t)slots + 1), *slotStart = x.template alloc<uint32_t>((size_t)slots + 1);
            x.run(slots, BeamSlots<false>{L, P, G, deepChunk, tri9, remap, slotBase, slotCnt, slotStart});
            L.poolSize = x.sum64(slotCnt, slots);
            totalPool += L.poolSize;
            if (totalPool >= (1ull << 32)) return "Subbeam pools exceed 2^32 entries: lower beam_max_depth (ptb_set_build_limits) or the grid size";
            x.scan(slotCnt, slotStart, slots + 1);
            L.pool = x.template alloc<uint32_t>((size_t)L.poolSize);
            x.run(slots, BeamSlots<true>{L, P, G, deepChunk, tri9, remap, slotBase, slotCnt, slotStart});
            x.run(L.n, BeamSlotGather{L, slotBase, slotStart});
        } else {
            x.run(L.n, BeamCount<false>{L, P, G, d, nTris, tri9, remap});
            L.poolSize = x.sum64(L.cnt, L.n);
            totalPool += L.poolSize; // bounds every u32 sum below (leaf entries are a subset of the level pools)
            if (totalPool >= (1ull << 32)) return "Subbeam pools exceed 2^32 entries: lower beam_max_depth (ptb_set_build_limits) or the grid size";
            x.scan(L.cnt, L.start, L.n);
            L.pool = x.template alloc<uint32_t>((size_t)L.poolSize);
            x.run(L.n, BeamCount<true>{L, P, G, d, nTris, tri9, remap});
        }
        x.run(L.n, BeamClassify{L, maxTris, d, maxDepth});
        const uint64_t nInner = x.sum64(L.inner, L.n);
        if (nInner == 0) break;
        if (16ull * nInner >= (1ull << 31) || d >= 30) return "Subbeam tree too large / too deep: set beam_max_depth (ptb_set_build_limits)";
        uint32_t *rank = x.template alloc<uint32_t>(L.n), *innerList = x.template alloc<uint32_t>((size_t)nInner);
        x.scan(L.inner, rank, L.n);
        x.run(L.n, BeamChildBase{L, rank, innerList});
        lv.push_back(new_level((uint32_t)(16ull * nInner)));
        x.run(lv[d + 1].n, BeamChildren{lv[d + 1], lv[d], innerList});
    }
    const int D = (int)lv.size() - 1;
    for (int d = D; d >= 0; --d) x.run(lv[d].n, BeamUp{lv[d], d < D ? lv[d + 1] : BeamLevelDev()});
    x.run(roots, BeamZeroBefore{lv[0]});
    for (int d = 1; d <= D; ++d) x.run(lv[d].n, BeamDown{lv[d], lv[d - 1]});
    uint32_t *desc = x.template alloc<uint32_t>(roots), *idxc = x.template alloc<uint32_t>(roots);
    uint32_t *descScan = x.template alloc<uint32_t>(roots), *idxScan = x.template alloc<uint32_t>(roots);
    x.run(roots, BeamRootTotals{lv[0], desc, idxc});
    const uint64_t nDesc = x.sum64(desc, roots), nIdx = x.sum64(idxc, roots);
    if (roots + nDesc >= (1ull << 32) || nIdx >= (1ull << 32)) return "Subbeam pools exceed 2^32 entries: lower beam_max_depth (ptb_set_build_limits) or the grid size";
    x.scan(desc, descScan, roots);
    x.scan(idxc, idxScan, roots);
    out.nBeams = roots + nDesc;
    out.nIdx = nIdx;
    out.depth = D;
    out.beams = x.template alloc_out<float>(12 * (size_t)out.nBeams);
    out.idx = x.template alloc_out<uint32_t>((size_t)nIdx);
    for (int d = 0; d <= D; ++d)
        x.run(lv[d].n, BeamEmit{lv[d], d > 0 ? lv[d - 1] : BeamLevelDev(), d, roots, descScan, idxScan, out.beams, out.idx});
    if (x.fail()) return "device error while building the beam levels";
    return "";
}

} // namespace ptb
