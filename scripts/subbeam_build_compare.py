#!/usr/bin/env python
"""Subbeam construction: the host's depth-first builder vs the level-synchronous builder on the GPU — build times and
byte equality of the pools (ptb_structure_digest).  Writes gpurun_out/subbeam_build.json."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import ptb_b200  # noqa: E402

sc = ptb_b200.scenes
cornell = sc.read_ptbs(os.path.join(ROOT, "tests", "golden", "cornel.ptbs"))
cases = [("Cornell, 202 triangles", cornell, (10, 40, 12)),
         ("Lucy stand-in, 524,310 triangles", sc.lucy_standin(cornell), (6, 50, 1)),
         ("soup 1e5", sc.triangle_soup(100_000), (4, 40, 2))]
out = {"host_cpus": os.cpu_count(), "cases": []}
for name, scene, (gs, m, depth) in cases:
    pt = ptb_b200.PathTracer(0)
    sc.upload_scene(pt, scene)
    pt.build_bvh()
    rec = {"scene": name, "params": f"gs {gs} / m {m} / depth guard {depth}"}
    wall, dev = pt.build_subbeams(gs, m, max_depth=depth, gpu=True)
    rec.update(gpu_wall_ms_first=wall, gpu_stream_ms_first=dev)
    wall, dev = pt.build_subbeams(gs, m, max_depth=depth, gpu=True)  # module loaded, allocator warm
    rec.update(gpu_wall_ms=wall, gpu_stream_ms=dev, gpu_digest=f"{pt.structure_digest(3):016x}", bytes=pt.structure_bytes(3))
    if "--no-host" not in sys.argv:
        t0 = time.perf_counter()
        pt.build_subbeams(gs, m, max_depth=depth, gpu=False)
        rec.update(host_wall_ms=(time.perf_counter() - t0) * 1e3, host_digest=f"{pt.structure_digest(3):016x}")
        rec["identical"] = rec["host_digest"] == rec["gpu_digest"]
    pt.close()
    print(json.dumps(rec), flush=True)
    out["cases"].append(rec)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", "subbeam_build.json"), "w"), indent=1)
