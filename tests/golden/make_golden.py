"""Generates the committed golden vectors from the UNMODIFIED reference compiled here
(oracle/_ref/ptref, built by `make -C oracle ref` from /root/reference/src).  Only runs in the
build container (the reference sources do not travel); the outputs do:

  tests/golden/cornel.ptbs            the reference's in-memory scene after initScene
                                      (scene.cpp:97-131) for showcase/scenes/cornel.scene
  tests/golden/cornel_ref_64.npz      per mode (ALL, BVH, TWO_PLANE, SUB_BEAMS), 64x64, seeds x<<16|y:
        t{m}, tri{m}, mat{m}          primary hit records (triangle ids through --wrap)
        rays                          the primary rays setupRay produced (tracer.cpp:72-88)
        img{m}_nee{0,1}               fp32 image after 2 spp of the unmodified trace(true)
        cnt{m}_nee{0,1}               [tri_tests, as_tests] (tracer.cpp:60-70)
        imgd{m}                       image with max_depth = 2, 3 spp, nee 1 (the cap is a --wrap of shade)

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle_lib import PTREF, ptref_primary_hits, ptref_render, run_ptref  # noqa: E402

REF_SHOWCASE = "/root/reference/showcase"
W = H = 64


def main():
    if not os.path.isfile(PTREF):
        raise SystemExit("oracle/_ref/ptref missing: make -C oracle ref (needs /root/reference)")
    ptbs = os.path.join(HERE, "cornel.ptbs")
    run_ptref(["--scene-file", "scenes/cornel.scene", "--base", REF_SHOWCASE, "dump-scene", ptbs])
    out = {}
    for m in range(4):
        t, tri, mat, rays, _ = ptref_primary_hits(ptbs, W, H, m)
        out[f"t{m}"], out[f"tri{m}"], out[f"mat{m}"] = t, tri, mat
        if m == 1:
            out["rays"] = rays
        for nee in (0, 1):
            img, rec, _ = ptref_render(ptbs, W, H, m, 2, 0, nee)
            out[f"img{m}_nee{nee}"] = img
            out[f"cnt{m}_nee{nee}"] = np.array([rec["tri_tests"], rec["as_tests"]], np.uint64)
        img, rec, _ = ptref_render(ptbs, W, H, m, 3, 2, 1)
        out[f"imgd{m}"] = img
    np.savez_compressed(os.path.join(HERE, "cornel_ref_64.npz"), **out)
    print("wrote", os.path.join(HERE, "cornel_ref_64.npz"))


if __name__ == "__main__":
    main()
