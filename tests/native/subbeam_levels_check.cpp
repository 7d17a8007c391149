// CPU check of the level-synchronous Subbeam builder (csrc/ptb_beam_levels.h): the driver template and its functors —
// the code ptb_beam_gpu.cu runs as CUDA kernels — executed with plain loops must reproduce build_subbeams (the depth-first
// host builder) byte for byte: beam records, leaf slices, their numbering.  Compiled and run by tests/test_native_host.py.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <vector>

#include "ptb_beam_levels.h"
#include "ptb_bvh_build.h"
#include "ptb_grid_build.h"
using namespace ptb;

struct HostExec {
    std::vector<void *> scratch;
    template <class T> T *alloc(size_t n) { void *p = calloc(n ? n : 1, sizeof(T)); scratch.push_back(p); return (T *)p; }
    template <class T> T *alloc_out(size_t n) { return (T *)calloc(n ? n : 1, sizeof(T)); }
    template <class F> void run(uint32_t n, const F &f) { for (uint32_t i = 0; i < n; ++i) f(i); }
    void scan(const uint32_t *in, uint32_t *out, uint32_t n) { uint32_t a = 0; for (uint32_t i = 0; i < n; ++i) { uint32_t v = in[i]; out[i] = a; a += v; } }
    uint64_t sum64(const uint32_t *in, uint32_t n) { uint64_t a = 0; for (uint32_t i = 0; i < n; ++i) a += in[i]; return a; }
    bool fail() { return false; }
    ~HostExec() { for (void *p : scratch) free(p); }
};

int main() {
    int bad = 0;
    struct Case { int n; uint32_t gs, m; int depth; float spread; };
    const Case cases[] = {{0, 3, 4, 2, 6}, {1, 2, 1, 3, 6}, {40, 2, 4, 3, 20}, {300, 3, 12, 2, 6}, {2000, 3, 40, 2, 3}, {5000, 2, 60, 2, 2}, {500, 2, 8, 3, 2}};
    for (const Case &cs : cases) {
        const int n = cs.n;
        std::mt19937 r(n + 7); std::uniform_real_distribution<float> U(-100, 100), S(-cs.spread, cs.spread);
        std::vector<float> pos(9 * n + 1);
        for (int i = 0; i < n; ++i) { float c[3] = {U(r), U(r), U(r)}; for (int v = 0; v < 3; ++v) for (int k = 0; k < 3; ++k) pos[9 * i + 3 * v + k] = c[k] + S(r); }
        BvhResult b; build_bvh(pos.data(), n, 3, b);
        std::vector<float> leaf(9 * n + 1);
        for (int k = 0; k < n; ++k) for (int j = 0; j < 9; ++j) leaf[9 * k + j] = pos[9 * b.order[k] + j];
        SubbeamResult R; build_subbeams(leaf.data(), n, b.lo, b.hi, cs.gs, cs.m, cs.depth, R);
        BeamGeom G;
        memcpy(G.gmin, R.gmin, sizeof G.gmin); memcpy(G.gmax, R.gmax, sizeof G.gmax);
        G.gs = cs.gs; G.cells = cs.gs * cs.gs * cs.gs * cs.gs; G.gsf = (float)cs.gs; G.inv = 1.0f / G.gsf;
        HostExec x; BeamLevelsResult out;
        // vertices by reference id + the leaf order as remap, like the device builder reads them
        const char *e = build_beam_levels(x, G, (uint32_t)n, pos.data(), b.order.data(), cs.m, cs.depth, out);
        bool ok = !*e && out.nBeams == R.beams.size() && out.nIdx == R.idx.size() &&
                  !memcmp(out.beams, R.beams.data(), R.beams.size() * sizeof(BeamRec)) && !memcmp(out.idx, R.idx.data(), R.idx.size() * 4);
        for (uint32_t deep : {1u, 64u}) { // the opt-in (beam, chunk-of-the-parent's-list) workers below the roots
            HostExec y; BeamLevelsResult o2;
            const char *e2 = build_beam_levels(y, G, (uint32_t)n, pos.data(), b.order.data(), cs.m, cs.depth, o2, deep);
            ok = ok && !*e2 && o2.nBeams == R.beams.size() && o2.nIdx == R.idx.size() &&
                 !memcmp(o2.beams, R.beams.data(), R.beams.size() * sizeof(BeamRec)) && !memcmp(o2.idx, R.idx.data(), R.idx.size() * 4);
            free(o2.beams); free(o2.idx);
        }
        printf("n=%d gs=%u m=%u depth=%d beams=%zu idx=%zu levels=%d %s %s\n", n, cs.gs, cs.m, cs.depth, R.beams.size(), R.idx.size(), out.depth, e, ok ? "IDENTICAL" : "DIFFERENT");
        if (!ok) bad = 1;
        free(out.beams); free(out.idx);
    }
    return bad;
}
