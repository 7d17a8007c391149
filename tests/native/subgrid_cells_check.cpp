// CPU check of the device Subgrid builder's per-channel routine (csrc/ptb_grid_gpu.h: populate_cell, shared between the
// kernels of ptb_grid_gpu.cu and the host): subgrid_topology + populate_cell over every (grid, channel), packed the way
// count -> scan -> fill packs them, must reproduce build_subgrid's pools byte for byte.  Compiled and run by
// tests/test_native_host.py with g++ (no GPU).  Prints one line per scene; exit code 1 on any difference.
#include "ptb_bvh_build.h"
#include "ptb_grid_build.h"
#include "ptb_grid_gpu.h"
#include <cstdio>
#include <cstring>
#include <random>
using namespace ptb;
int main(){
  int bad = 0;
  for (int n : {0, 1, 5, 200, 5000}) {
    std::mt19937 r(n); std::uniform_real_distribution<float> U(-100,100), S(-6,6);
    std::vector<float> pos(9*n+1);
    for(int i=0;i<n;++i){float c[3]={U(r),U(r),U(r)};for(int v=0;v<3;++v)for(int k=0;k<3;++k)pos[9*i+3*v+k]=c[k]+S(r);}
    BvhResult b; build_bvh(pos.data(), n, 3, b);
    std::vector<float> leaf(9*n+1); for(int k=0;k<n;++k) for(int j=0;j<9;++j) leaf[9*k+j]=pos[9*b.order[k]+j];
    SubgridResult R; build_subgrid(leaf.data(), n, b, 6, 4, 40, R);
    SubgridTopo T; subgrid_topology(n, b, 6, 4, 40, T);
    bool ok = T.grids.size()==R.hdr.size() && T.cell_base.back()==R.lut.size();
    std::vector<GridCell> lut(T.cell_base.back()); std::vector<uint32_t> idx;
    std::vector<uint32_t> tmp(T.max_cand+1);
    for (size_t g=0; g<T.grids.size() && ok; ++g) {
      GridHdr h = subgrid_header(T.grids[g], T.cell_base[g]);
      if (memcmp(&h, &R.hdr[g], 64)) { ok=false; printf("hdr %zu differs\n", g); }
      for (uint64_t c=T.cell_base[g]; c<T.cell_base[g+1]; ++c) {
        unsigned long long mask;
        // emulate remap: vtx by reference id + id_of_tri
        uint32_t cnt = populate_cell(T.grids[g], (uint32_t)(c-T.cell_base[g]), T.grids.data(), T.cand.data(), pos.data(), b.order.data(), &mask, tmp.data());
        lut[c] = {(uint32_t)idx.size(), (uint32_t)idx.size()+cnt};
        idx.insert(idx.end(), tmp.begin(), tmp.begin()+cnt);
        if (T.grids[g].cand_count <= 64 && (uint32_t)__builtin_popcountll(mask) != cnt) ok=false;
      }
    }
    ok = ok && idx.size()==R.idx.size() && !memcmp(idx.data(), R.idx.data(), idx.size()*4) && !memcmp(lut.data(), R.lut.data(), lut.size()*8);
    printf("n=%d grids=%zu cells=%zu idx=%zu max_cand=%u %s\n", n, T.grids.size(), lut.size(), idx.size(), T.max_cand, ok?"IDENTICAL":"DIFFERENT");
    if (!ok) bad = 1;
  }
  return bad;
}
