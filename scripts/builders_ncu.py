#!/usr/bin/env python
"""The structure builders under ncu: one BVH build on the GPU (soup of N triangles) and one Subgrid build (Lucy stand-in
level L), nothing else — for
  ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv \
      --log-file gpurun_out/builders_ncu.csv python scripts/builders_ncu.py [log10_n=7] [lucy_level=7]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import ptb_b200  # noqa: E402

sc = ptb_b200.scenes
lg = int(sys.argv[1]) if len(sys.argv) > 1 else 7
level = int(sys.argv[2]) if len(sys.argv) > 2 else 7
pt = ptb_b200.PathTracer(0)
sc.upload_scene(pt, sc.triangle_soup(10 ** lg))
print("bvh_gpu_ms", pt.build_bvh(gpu=True), "depth", pt.bvh_depth())
pt.close()
pt = ptb_b200.PathTracer(0)
sc.upload_scene(pt, sc.lucy_standin(sc.read_ptbs(os.path.join(ROOT, "tests", "golden", "cornel.ptbs")), level=level))
pt.build_bvh(gpu=True)
print("subgrid_gpu_ms", pt.build_subgrid(8, 8, 50, gpu=True), "bytes", pt.structure_bytes(2))
pt.close()
