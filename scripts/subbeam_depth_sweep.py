#!/usr/bin/env python
"""Subbeams on the Lucy stand-in: what does the depth guard cost and buy?  For each beam_max_depth: GPU build time, structure
bytes, primary-ray Mrays/s (1080p) and full-path Mrays/s at 1 spp.  Stops before a build that would not fit (the structure
grows ~16x per level until beams thin out).  usage (under gpurun): scripts/subbeam_depth_sweep.py [gs] [m] [depths...]"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import ptb_b200  # noqa: E402

gs = int(sys.argv[1]) if len(sys.argv) > 1 else 6
m = int(sys.argv[2]) if len(sys.argv) > 2 else 50
depths = [int(a) for a in sys.argv[3:]] or [1, 2, 3]
sc = ptb_b200.scenes
pt = ptb_b200.PathTracer(0)
sc.upload_scene(pt, sc.lucy_standin(sc.read_ptbs(os.path.join(ROOT, "tests/golden/cornel.ptbs"))))
pt.build_bvh()
W, H = 1920, 1080
pt.set_params(width=W, height=H, mode=1, nee=0, max_depth=9)
t_bvh, tri_bvh, _ = pt.primary_hits()
last_bytes = 0
for d in depths:
    if last_bytes * 16 > 60e9:
        print(json.dumps({"gs": gs, "m": m, "beam_max_depth": d, "skipped": "previous level x16 would exceed 60 GB"}))
        break
    try:
        t0 = time.perf_counter()
        wall, dev = pt.build_subbeams(gs, m, max_depth=d, gpu=True)
    except ptb_b200.PtbError as e:
        print(json.dumps({"gs": gs, "m": m, "beam_max_depth": d, "error": str(e)[:200]}))
        break
    nbytes = pt.structure_bytes(3)
    last_bytes = nbytes
    pt.set_params(width=W, height=H, mode=3, nee=0, max_depth=9)
    t0 = time.perf_counter()
    t, tri, _ = pt.primary_hits()
    prim_s = time.perf_counter() - t0
    pt.reset()
    pt.render(1)
    ms = pt.last_ms()
    rays = pt.counters()["rays"]
    print(json.dumps({"gs": gs, "m": m, "beam_max_depth": d, "build_wall_ms": wall, "build_device_ms": dev, "bytes": nbytes,
                      "primary_mrays_s_incl_copies": W * H / prim_s / 1e6, "path_1spp_ms": ms, "path_mrays_s": rays / ms / 1e3,
                      "primary_t_equal_bvh_frac": float((t.view(np.uint32) == t_bvh.view(np.uint32)).mean())}), flush=True)
pt.close()
