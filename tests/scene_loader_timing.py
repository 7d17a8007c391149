#!/usr/bin/env python
"""Scene-text loading time, this repo's loader (ptb_convert_scene_file, csrc/ptb_scene_loader.cpp) beside the reference's
own loadScene + loadObject as compiled into oracle/_ref/ptref, on a generated OBJ of N triangles (default 1e6; v/vt/vn
corners, one material), and byte equality of the two .ptbs dumps.  Lives under tests/ because it executes the reference
binary.  Usage: python tests/scene_loader_timing.py [N] -> profiles/r1_scene_loader_timing.json"""
import json
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import ptb_b200  # noqa: E402
from oracle_lib import have_ref, run_ptref  # noqa: E402

n = int(float(sys.argv[1])) if len(sys.argv) > 1 else 1_000_000
rng = np.random.default_rng(3)
d = tempfile.mkdtemp()
os.makedirs(os.path.join(d, "objs"))
nv = n // 2 + 2
v = rng.uniform(-1, 1, (nv, 3))
vn = rng.normal(size=(1000, 3))
vn /= np.linalg.norm(vn, axis=1, keepdims=True)
vt = rng.uniform(0, 1, (1000, 2))
f = rng.integers(1, nv + 1, (n, 3))
k = rng.integers(1, 1001, (n, 6))
with open(os.path.join(d, "objs", "big.obj"), "w") as o:
    o.write("mtllib big.mtl\no big\n")
    o.write("".join("v %.6f %.6f %.6f\n" % tuple(r) for r in v))
    o.write("".join("vn %.4f %.4f %.4f\n" % tuple(r) for r in vn))
    o.write("".join("vt %.6f %.6f\n" % tuple(r) for r in vt))
    o.write("usemtl a\n")
    o.write("".join("f %d/%d/%d %d/%d/%d %d/%d/%d\n" % (a, k0, k1, b, k2, k3, c, k4, k5) for (a, b, c), (k0, k1, k2, k3, k4, k5) in zip(f, k)))
with open(os.path.join(d, "objs", "big.mtl"), "w") as o:
    o.write("newmtl a\nKd 0.5 0.6 0.7\n")
scene = os.path.join(d, "big.scene")
with open(scene, "w") as o:
    o.write("[Models]\nfile: objs/big.obj\npos: 0.5,-2,3\nscale: 100,90,-110\n\n[Camera]\npos: 0,0,-500\nforward: 0,0,1\nfocus: 1\n")
res = {"triangles": n, "obj_bytes": os.path.getsize(os.path.join(d, "objs", "big.obj")), "host_cpus": os.cpu_count()}
best = 1e30
for _ in range(3):
    t0 = time.perf_counter()
    ptb_b200.convert_scene_file(scene, os.path.join(d, "mine.ptbs"), base_dir=d)
    best = min(best, time.perf_counter() - t0)
res["this_repo_load_and_write_s"] = best
if have_ref():
    rec = [r for r in run_ptref(["--scene-file", scene, "--base", d, "dump-scene", os.path.join(d, "ref.ptbs")]) if r["cmd"] == "dump-scene"][0]
    res["reference_load_s"] = rec["load_ms"] / 1e3
    with open(os.path.join(d, "mine.ptbs"), "rb") as a, open(os.path.join(d, "ref.ptbs"), "rb") as b:
        res["identical"] = a.read() == b.read()
print(json.dumps(res))
json.dump(res, open(os.path.join(ROOT, "profiles", "r1_scene_loader_timing.json"), "w"), indent=1)
