#!/usr/bin/env python
"""BASELINE config 3: Lucy stand-in, 1920x1080 primary rays, traversal-only throughput per structure and
hit-id diff against the reference (tests/golden/lucy_standin_ref_1080p.npz).  Writes gpurun_out/config3.json."""
import hashlib
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import ptb_b200  # noqa: E402

sc = ptb_b200.scenes
scene = sc.lucy_standin(sc.read_ptbs(os.path.join(ROOT, "tests", "golden", "cornel.ptbs")))
gold_path = os.path.join(ROOT, "tests", "golden", "lucy_standin_ref_1080p.npz")
gold = np.load(gold_path) if os.path.isfile(gold_path) else None
W, H = 1920, 1080
pt = ptb_b200.PathTracer(0)
sc.upload_scene(pt, scene)
out = {"triangles": int(scene["pos"].shape[0]), "width": W, "height": H, "structures": {}}
if gold is not None:
    out["scene_matches_golden"] = bool(hashlib.sha256(scene["pos"].tobytes()).hexdigest() == str(gold["scene_sha256"]))
    out["reference_cpu"] = {"bvh_ms": float(gold["ref_bvh_ms"]), "threads": int(gold["ref_threads"]),
                            "bvh_mrays_s": W * H / float(gold["ref_bvh_ms"]) / 1e3}
    if "ref_two_plane_ms" in gold:
        out["reference_cpu"].update(two_plane_ms=float(gold["ref_two_plane_ms"]), two_plane_build_s=float(gold["ref_two_plane_build"]) / 1e3,
                                    two_plane_bytes=int(gold["ref_two_plane_bytes"]),
                                    two_plane_mrays_s=W * H / float(gold["ref_two_plane_ms"]) / 1e3)
builds = [("BVH", 1, lambda: pt.build_bvh()), ("Subgrid gs8 os8 m50", 2, lambda: pt.build_subgrid(8, 8, 50, gpu=True)),
          ("Subbeams gs6 m50 depth2", 3, lambda: pt.build_subbeams(6, 50, max_depth=2, gpu=True))]
t_ref = None
for name, mode, build in builds:
    t0 = time.perf_counter()
    build()
    bs = time.perf_counter() - t0
    pt.set_params(width=W, height=H, mode=mode, count_tests=0)
    pt.primary_hits()
    ms = []
    for _ in range(5):
        t, tri, mat = pt.primary_hits()
        ms.append(pt.last_ms())
    rec = {"build_s": bs, "bytes": pt.structure_bytes(mode), "ms": min(ms), "mrays_s": W * H / min(ms) / 1e3,
           "hit_fraction": float((t < np.inf).mean())}
    if mode == 1:
        t_ref, tri_ref = t, tri
    else:
        rec["t_bit_equal_to_bvh"] = bool((t.view(np.uint32) == t_ref.view(np.uint32)).all())
        rec["id_agreement_with_bvh"] = float((tri == tri_ref).mean())
    if gold is not None and out["scene_matches_golden"]:
        rec["t_digest_equals_reference"] = bool(hashlib.sha256(t.tobytes()).hexdigest() == str(gold["t_sha256"]))
        rec["id_agreement_with_reference"] = float((tri == gold["tri"]).mean())
        rec["id_mismatches"] = int((tri != gold["tri"]).sum())
    out["structures"][name] = rec
    print(name, rec, flush=True)
pt.close()
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "config3.json"), "w"), indent=1)
