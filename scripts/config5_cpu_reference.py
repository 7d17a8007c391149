#!/usr/bin/env python
"""The CPU-reference side of BASELINE config 5 (capped by the reference's builders — BVH build 7 s at 1e5 triangles and
superlinear, Subgrid build 64 s at 1e5): the UNMODIFIED reference (oracle/_ref/ptref trace-rays: findIntersection from all host
threads over a caller-supplied batch) on the same soups and the same incoherent rays as scripts/config5_soup_sweep.py, 2^20 rays.
CPU only; run where /root/reference was compiled.  usage: scripts/config5_cpu_reference.py [max_log10_n=5] > profiles/...jsonl"""
import json
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import ptb_b200  # noqa: E402
from oracle_lib import ptref_trace_rays  # noqa: E402

sc = ptb_b200.scenes
max_log = int(sys.argv[1]) if len(sys.argv) > 1 else 5
rays = sc.soup_rays(1 << 20)
with tempfile.TemporaryDirectory() as d:
    for lg in range(4, max_log + 1):
        n = 10 ** lg
        path = os.path.join(d, f"soup{lg}.ptbs")
        sc.write_ptbs(path, sc.triangle_soup(n))
        for name, mode, extra in (("BVH", 1, ()), ("Subgrid gs6 os4 m40", 2, ("--gs", 6, "--os", 4, "--m", 40))):
            if mode == 2 and lg > 4:
                continue  # the reference's constructGrid needs 64 s and 377 MB at 1e5 triangles already
            t, tri, recs = ptref_trace_rays(path, rays, mode, extra=extra)
            tr = [r for r in recs if r["cmd"] == "trace-rays"][0]
            b = [r for r in recs if r["cmd"] == "build"][0]
            print(json.dumps({"n": n, "structure": name, "rays": tr["rays"], "threads": tr["threads"], "host_cpus": os.cpu_count(),
                              "mrays_s": tr["mrays_s"], "trace_ms": tr["ms"], "bvh_build_ms": b["bvh_ms"], "grid_build_ms": b["grid_ms"],
                              "hit_fraction": float((t < np.inf).mean())}), flush=True)
