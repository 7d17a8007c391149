#!/usr/bin/env python
"""Subbeam builder on the GPU: one thread per beam below the roots (default) against the chunked workers
(PTB_BEAM_DEEP_CHUNK=k: one thread per (beam, k entries of the parent's list)) on the Lucy stand-in, gs 6 / m 50.
Each setting in its own process (the knob is read from the environment); prints build time and the pools' digest.
usage (under gpurun): scripts/subbeam_deep_chunk.py [depth=2] [k ...]"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
depth = int(sys.argv[1]) if len(sys.argv) > 1 else 2
ks = sys.argv[2:] or ["0", "64", "256", "1024"]
CHILD = r'''
import json, os, sys
sys.path.insert(0, %r)
import ptb_b200
sc = ptb_b200.scenes
pt = ptb_b200.PathTracer(0)
sc.upload_scene(pt, sc.lucy_standin(sc.read_ptbs(os.path.join(%r, "tests/golden/cornel.ptbs"))))
pt.build_bvh()
best = None
for _ in range(2):
    wall, dev = pt.build_subbeams(6, 50, max_depth=%d, gpu=True)
    best = (wall, dev) if best is None or dev < best[1] else best
print("REC " + json.dumps({"wall_ms": best[0], "device_ms": best[1], "digest": "%%016x" %% pt.structure_digest(3), "bytes": pt.structure_bytes(3)}))
''' % (ROOT, ROOT, depth)
for k in ks:
    env = dict(os.environ)
    env.pop("PTB_BEAM_DEEP_CHUNK", None)
    if k != "0":
        env["PTB_BEAM_DEEP_CHUNK"] = k
    r = subprocess.run([sys.executable, "-c", CHILD], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600)
    rec = [l for l in r.stdout.splitlines() if l.startswith("REC ")]
    print(json.dumps(dict(depth=depth, deep_chunk=int(k), **(json.loads(rec[0][4:]) if rec else {"failed": r.stderr[-300:]}))), flush=True)
