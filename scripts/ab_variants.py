#!/usr/bin/env python
"""A/B of libptb variants on one GPU (run under gpurun): for each library — the default build and every
bachelor-thesis-pathtracer_b200/variants/libptb_*.so, or those named on the command line — a child process checks the frames
against the reference goldens (tests/golden/configs_ref.npz: Cornell 1024^2 x 4 spp, Lucy stand-in 1080p x 2 spp, by digest)
and times full path tracing (device time of ptb_render, best of `reps`): Lucy BVH 16 spp, Cornell BVH 64 spp, Lucy Subgrid 4 spp.
usage: scripts/ab_variants.py [--no-subgrid] [--reps 3] [name ...] > gpurun_out/ab.txt"""
import glob
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r'''
import hashlib, json, os, sys
sys.path.insert(0, %(root)r); sys.path.insert(0, os.path.join(%(root)r, "tests"))
import numpy as np, ptb_b200
sc = ptb_b200.scenes
gold = np.load(os.path.join(%(root)r, "tests/golden/configs_ref.npz"))
cornel = os.path.join(%(root)r, "tests/golden/cornel.ptbs")
reps, subgrid = %(reps)d, %(subgrid)d
out = {}
def run(pt, W, H, spp, depth, mode, reps):
    best, rays = 1e30, 0
    for _ in range(reps):
        pt.set_params(width=W, height=H, mode=mode, nee=0, max_depth=depth)
        pt.reset(); pt.render(spp)
        best = min(best, pt.last_ms()); rays = pt.counters()["rays"]
    return rays / best / 1e3
pt = ptb_b200.PathTracer(0); pt.load_ptbs(cornel); pt.build_bvh()
pt.set_params(width=1024, height=1024, mode=1, nee=0, max_depth=5); pt.reset(); pt.render(4)
out["c1_equal"] = hashlib.sha256(pt.read_image().tobytes()).hexdigest() == str(gold["c1_img_sha256"])
out["cornell"] = run(pt, 1024, 1024, 64, 5, 1, reps)
pt.close()
pt = ptb_b200.PathTracer(0); sc.upload_scene(pt, sc.lucy_standin(sc.read_ptbs(cornel))); pt.build_bvh()
pt.set_params(width=1920, height=1080, mode=1, nee=0, max_depth=9); pt.reset(); pt.render(2)
img = pt.read_image()
out["lucy_equal"] = hashlib.sha256(img.tobytes()).hexdigest() == str(gold["lucy_bvh_img_sha256"])
out["lucy_digest"] = hashlib.sha256(img.tobytes()).hexdigest()[:12]
out["lucy"] = run(pt, 1920, 1080, 16, 9, 1, reps)
if subgrid:
    pt.build_subgrid(8, 8, 50, gpu=True)
    pt.set_params(width=1920, height=1080, mode=2, nee=0, max_depth=9); pt.reset(); pt.render(2)
    out["grid_digest"] = hashlib.sha256(pt.read_image().tobytes()).hexdigest()[:12]
    out["lucy_subgrid"] = run(pt, 1920, 1080, 4, 9, 2, max(1, reps - 1))
pt.close()
print("AB_JSON " + json.dumps(out))
'''


def main():
    args = sys.argv[1:]
    subgrid = 0 if "--no-subgrid" in args else 1
    reps = int(args[args.index("--reps") + 1]) if "--reps" in args else 3
    names = [a for i, a in enumerate(args) if not a.startswith("--") and (i == 0 or args[i - 1] != "--reps")]
    vdir = os.path.join(ROOT, "bachelor-thesis-pathtracer_b200", "variants")
    libs = [("default", "")] + [(os.path.basename(p)[7:-3], p) for p in sorted(glob.glob(os.path.join(vdir, "libptb_*.so")))]
    if names:
        libs = [l for l in libs if l[0] in names]
    env_extra = {k: v for k, v in os.environ.items() if k.startswith("PTB_") and k != "PTB_LIB"}
    print(f"# env {env_extra}  reps {reps}", flush=True)
    for name, path in libs:
        env = dict(os.environ)
        if path:
            env["PTB_LIB"] = path
        r = subprocess.run([sys.executable, "-c", CHILD % dict(root=ROOT, reps=reps, subgrid=subgrid)], env=env, stdout=subprocess.PIPE,
                           stderr=subprocess.PIPE, text=True, timeout=900)
        rec = [l for l in r.stdout.splitlines() if l.startswith("AB_JSON ")]
        if r.returncode != 0 or not rec:
            print(f"{name:16s} FAILED rc={r.returncode} {r.stderr[-300:]!r}", flush=True)
            continue
        d = json.loads(rec[0][8:])
        print(f"{name:16s} lucy {d['lucy']:7.0f}  cornell {d['cornell']:7.0f}  lucy_subgrid {d.get('lucy_subgrid', 0):7.0f}  Mrays/s   "
              f"c1 digest {'==' if d['c1_equal'] else 'DIFFERS'}  lucy digest {'==' if d['lucy_equal'] else d['lucy_digest']}  grid {d.get('grid_digest', '-')}",
              flush=True)


if __name__ == "__main__":
    main()
