#!/usr/bin/env python
"""Where does the Lucy stand-in's 1080p BVH frame differ from the reference's (1 pixel in 2,073,600 at 2 spp)?
Finds the pixel(s), tells whether the first or the second sample diverged and whether the RNG stream did, and — to tell a
rounding of the hand-rolled sincos from a wrong hit — compares 2^22 incoherent rays through the scene hit for hit with the
oracle (intersect_tri's conservative pre-rejects must never drop a hit the reference's order accepts).
Run under gpurun; uses the oracle as the checker (test infrastructure)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import ptb_b200  # noqa: E402
from oracle_lib import Oracle  # noqa: E402

sc = ptb_b200.scenes
scene = sc.lucy_standin(sc.read_ptbs(os.path.join(ROOT, "tests/golden/cornel.ptbs")))
path = "/tmp/lucy_diag.ptbs"
sc.write_ptbs(path, scene)
W, H, depth = 1920, 1080, 9
o = Oracle(path)
o.build()
pt = ptb_b200.PathTracer(0)
pt.load_ptbs(path)
pt.build_bvh()
pt.set_params(width=W, height=H, mode=1, nee=0, max_depth=depth)
o.configure(W, H, 1, nee=0, max_depth=depth)
res = {}
for spp in (1, 2, 4):
    pt.reset()
    pt.render(spp)
    g_img, (g_s, g_rng) = pt.read_image().copy(), pt.read_state()
    r_img, r_s, r_rng, _ = o.render(spp, threads=os.cpu_count())
    diff = (g_img.view(np.uint32) != r_img.view(np.uint32)).any(axis=-1)
    rdiff = (g_rng != r_rng).reshape(H, W)
    print(f"spp {spp}: {int(diff.sum())} pixels differ, {int(rdiff.sum())} RNG end states differ")
    for y, x in zip(*np.nonzero(diff | rdiff)):
        print(f"   pixel ({x},{y}): gpu {g_img[y, x, :3]} ref {r_img[y, x, :3]}  rng gpu {g_rng[y * W + x]:#x} ref {r_rng[y * W + x]:#x}")
# hit-for-hit on incoherent rays
rng = np.random.default_rng(99)
n = 1 << 22
org = rng.normal(size=(n, 3))
org = 600.0 * org / np.linalg.norm(org, axis=1, keepdims=True)
tgt = rng.uniform(-250, 250, size=(n, 3))
tgt[: n // 2] = rng.uniform(-120, 120, size=(n // 2, 3)) + np.array([0.0, -100.0, 0.0])  # half of them into the statue
d = tgt - org
d /= np.linalg.norm(d, axis=1, keepdims=True)
od = np.concatenate([org, d], axis=1).astype(np.float32)
gt, gtri = pt.trace_rays(od, 1)
rt, rtri = o.trace_rays(od, threads=os.cpu_count())
bad_t = int((gt.view(np.uint32) != rt.view(np.uint32)).sum())
bad_id = int(((gtri != rtri) & (gt.view(np.uint32) == rt.view(np.uint32))).sum())
print(f"{n} incoherent rays: {bad_t} hit distances differ, {bad_id} ids differ at equal distance (ties); hit fraction {(gt < np.inf).mean():.3f}")
for i in np.nonzero(gt.view(np.uint32) != rt.view(np.uint32))[0][:5]:
    print(f"   ray {i}: gpu t {gt[i]!r} tri {gtri[i]}  ref t {rt[i]!r} tri {rtri[i]}  od {od[i]}")
