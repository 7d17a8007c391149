#!/usr/bin/env python
"""Headless stand-in for `pathtracer <scene> -S n -gs g -os o -m m --testing [--2plane_only]` (main.cpp:36-141): the same
flags and defaults, the same five stdout lines per structure, frames under ./render/ with the reference's file names.
<scene> is a .scene file of the reference (with --base), a .ptbs file (tests/golden/cornel.ptbs; .ptbs is what oracle/ref/ref_driver.cpp dumps from the reference's
own loader) or one of the built-in names `cornell` / `lucy` (the 524,310-triangle stand-in of BASELINE config 3).
Example: python scripts/ptb_testing.py lucy -S 10 -gs 8 -os 8 -m 50 --testing | tail -n 10   (showcase/test.py:48)"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import ptb_b200  # noqa: E402

ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
ap.add_argument("scene")
ap.add_argument("-S", type=int, default=100, dest="samples")
ap.add_argument("-gs", type=int, default=10)
ap.add_argument("-os", type=int, default=5, dest="os_")
ap.add_argument("-m", type=int, default=40)
ap.add_argument("-ds", type=int, default=10, help="debug scale of the heat maps (main.cpp:62-67)")
ap.add_argument("--testing", action="store_true", help="accepted for command-line compatibility; this driver always runs the protocol")
ap.add_argument("--2plane_only", action="store_true", dest="two_plane_only")
ap.add_argument("--width", type=int, default=1280)   # setWindowSize(1280, 720), main.cpp:92
ap.add_argument("--height", type=int, default=720)
ap.add_argument("--out", default="./render")
ap.add_argument("--base", default="", help="directory the model paths of a .scene file are relative to")
ap.add_argument("--device", type=int, default=0)
a = ap.parse_args()

sc = ptb_b200.scenes
cornell = os.path.join(ROOT, "tests", "golden", "cornel.ptbs")
pt = ptb_b200.PathTracer(a.device)
if a.scene == "cornell":
    pt.load_ptbs(cornell)
elif a.scene == "lucy":
    sc.upload_scene(pt, sc.lucy_standin(sc.read_ptbs(cornell)))
elif a.scene.endswith(".scene"):  # the reference's own scene text; model paths relative to --base (the reference runs inside showcase/)
    pt.load_scene_file(a.scene, a.base, ignore_textures=True)
else:
    pt.load_ptbs(a.scene)
ptb_b200.testing_protocol.run(pt, samples=a.samples, gs=a.gs, os_=a.os_, m=a.m, two_plane_only=a.two_plane_only, width=a.width,
                              height=a.height, debug_scale=a.ds, out_dir=a.out)
pt.close()
