#!/usr/bin/env python
"""One GPU's share of a tile-sharded N-GPU frame, on one GPU: params rank = 0, world = N renders exactly the pixels rank 0 of N
would own.  Sweeps the resident CTAs per SM of the path kernel (PTB_MEGA_CTAS) and the samples per work item (PTB_MEGA_CHUNK)
— a pixel's samples are serial, so with ~1 pixel per lane the frame time is the latency of the slowest pixel chain, and fewer,
faster lanes may beat a full machine.  Each setting in its own process (the knobs are read at ptb_create).
usage (under gpurun): scripts/tile_share_sweep.py [world=8] [spp=256] [structure=bvh|subgrid]"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
world = int(sys.argv[1]) if len(sys.argv) > 1 else 8
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 256
structure = sys.argv[3] if len(sys.argv) > 3 else "bvh"
CHILD = r'''
import json, os, sys
sys.path.insert(0, %r)
import ptb_b200
sc = ptb_b200.scenes
pt = ptb_b200.PathTracer(0)
sc.upload_scene(pt, sc.lucy_standin(sc.read_ptbs(os.path.join(%r, "tests/golden/cornel.ptbs"))))
pt.build_bvh()
mode = 1
if %r == "subgrid":
    pt.build_subgrid(8, 8, 50, gpu=True); mode = 2
pt.set_params(width=1920, height=1080, mode=mode, nee=0, max_depth=9, rank=0, world=%d, tile_w=32, tile_h=8)
best = 1e30
for _ in range(3):
    pt.reset(); pt.render(%d)
    best = min(best, pt.last_ms()); rays = pt.counters()["rays"]
print("REC " + json.dumps({"ms": best, "mrays_s": rays / best / 1e3, "rays": rays}))
''' % (ROOT, ROOT, structure, world, spp)
full = None
for ctas in (12, 10, 8, 6, 4, 3):
    for chunk in (0, 1, 4, 16, 64):
        env = dict(os.environ, PTB_MEGA_CTAS=str(ctas))
        if chunk:
            env["PTB_MEGA_CHUNK"] = str(chunk)
        r = subprocess.run([sys.executable, "-c", CHILD], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600)
        rec = [l for l in r.stdout.splitlines() if l.startswith("REC ")]
        d = json.loads(rec[0][4:]) if rec else {"failed": r.stderr[-200:]}
        print(json.dumps(dict(world=world, spp=spp, structure=structure, ctas_per_sm=ctas, chunk=chunk or "auto", **d)), flush=True)
