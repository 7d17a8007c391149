#!/usr/bin/env python
"""Subgrid construction at BASELINE config 3's size (Lucy stand-in, gs 8 / os 8 / m 50): host builder vs the channel
population on the GPU — build times and byte equality of the pools.  Writes gpurun_out/subgrid_build.json."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import ptb_b200  # noqa: E402

sc = ptb_b200.scenes
scene = sc.lucy_standin(sc.read_ptbs(os.path.join(ROOT, "tests", "golden", "cornel.ptbs")))
pt = ptb_b200.PathTracer(0)
sc.upload_scene(pt, scene)
pt.build_bvh()
res = {"scene": "Lucy stand-in, 524,310 triangles", "params": "gs 8 / os 8 / m 50", "host_cpus": os.cpu_count()}
wall, dev = pt.build_subgrid(8, 8, 50, gpu=True)
res.update(gpu_wall_ms=wall, gpu_device_ms=dev, gpu_digest=f"{pt.structure_digest(2):016x}", bytes=pt.structure_bytes(2))
wall, dev = pt.build_subgrid(8, 8, 50, gpu=True)  # second build: allocator warm
res.update(gpu_wall_ms_2nd=wall, gpu_device_ms_2nd=dev)
if "--no-host" not in sys.argv:
    t0 = time.perf_counter()
    pt.build_subgrid(8, 8, 50, gpu=False)
    res.update(host_wall_ms=(time.perf_counter() - t0) * 1e3, host_digest=f"{pt.structure_digest(2):016x}")
    res["identical"] = res["host_digest"] == res["gpu_digest"]
pt.close()
print(json.dumps(res))
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(res, open(os.path.join(ROOT, "gpurun_out", "subgrid_build.json"), "w"), indent=1)
