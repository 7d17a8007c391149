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

  tests/golden/lucy_standin_ref_1080p.npz   (--lucy; the reference needs 25 s for its BVH, 16 min / 11 GB
                                      for its Subgrids (gs 8, os 8, m 50) on the 524,310-triangle stand-in)
        scene_sha256                  digest of the generated positions (the generator is bit-reproducible)
        tri                           primary hit triangle ids, 1920x1080, reference BVH mode
        t_sha256                      digest of the hit distances (bit pattern)
        two_plane_equal               1 if the reference's TWO_PLANE run gave identical t and ids

  tests/golden/cornel_textured_ref_48.npz   (--textured) the reference's texture branches (shader.cpp:52-66,251-258,
                                      texture.cpp:22-32) on scenes.textured_cornell: 64x48, 3 spp, BVH and TWO_PLANE
        img_m{mode}_nee{0,1}_d{0,4}_l{0,1}   fp32 image (l = the emissive cube carries a texture too)
        scene_sha256_l{0,1}           digest of the generated textures + uv (the generator is bit-reproducible)

    python tests/golden/make_golden.py [--lucy [--lucy-two-plane]] [--textured]
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle_lib import PTREF, ptref_primary_hits, ptref_render, run_ptref  # noqa: E402

REF_SHOWCASE = "/root/reference/showcase"
W = H = 64


def lucy():
    import hashlib
    import tempfile
    sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
    import ptb_b200
    sc = ptb_b200.scenes
    scene = sc.lucy_standin(sc.read_ptbs(os.path.join(HERE, "cornel.ptbs")))
    with tempfile.TemporaryDirectory() as d:
        ptbs = os.path.join(d, "lucy.ptbs")
        sc.write_ptbs(ptbs, scene)
        t, tri, mat, rays, recs = ptref_primary_hits(ptbs, 1920, 1080, 1)
        out = {"scene_sha256": hashlib.sha256(scene["pos"].tobytes()).hexdigest(), "tri": tri,
               "t_sha256": hashlib.sha256(t.tobytes()).hexdigest(), "hit_fraction": float((t < np.inf).mean()),
               "ref_bvh_ms": [r for r in recs if r["cmd"] == "primary-hits"][0]["ms"],
               "ref_threads": [r for r in recs if r["cmd"] == "primary-hits"][0]["threads"], "two_plane_equal": -1}
        if "--lucy-two-plane" in sys.argv:
            t2, tri2, _, _, recs2 = ptref_primary_hits(ptbs, 1920, 1080, 2, extra=["--gs", 8, "--os", 8, "--m", 50])
            out["two_plane_equal"] = int((t2.view(np.uint32) == t.view(np.uint32)).all() and (tri2 == tri).all())
            out["ref_two_plane_ms"] = [r for r in recs2 if r["cmd"] == "primary-hits"][0]["ms"]
            out["ref_two_plane_build"] = [r for r in recs2 if r["cmd"] == "build"][0]["grid_ms"]
            out["ref_two_plane_bytes"] = [r for r in recs2 if r["cmd"] == "build"][0]["grid_bytes"]
    np.savez_compressed(os.path.join(HERE, "lucy_standin_ref_1080p.npz"), **out)
    print("wrote lucy_standin_ref_1080p.npz", {k: v for k, v in out.items() if k != "tri"})


def textured_scene_digest(scene):
    import hashlib
    h = hashlib.sha256(scene["uv"].tobytes())
    for t in scene["textures"]:
        h.update(np.ascontiguousarray(t, np.float32).tobytes())
    return h.hexdigest()


def textured():
    import tempfile
    sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
    import ptb_b200
    sc = ptb_b200.scenes
    room = sc.read_ptbs(os.path.join(HERE, "cornel.ptbs"))
    out = {}
    with tempfile.TemporaryDirectory() as d:
        for lt in (0, 1):
            scene = sc.textured_cornell(room, light_texture=bool(lt))
            ptbs = os.path.join(d, f"tex{lt}.ptbs")
            sc.write_ptbs(ptbs, scene)
            out[f"scene_sha256_l{lt}"] = textured_scene_digest(scene)
            for mode in (1, 2):
                for nee, depth in ((0, 0), (1, 0), (1, 4)):
                    img, rec, _ = ptref_render(ptbs, 64, 48, mode, 3, depth, nee)
                    out[f"img_m{mode}_nee{nee}_d{depth}_l{lt}"] = img
    np.savez_compressed(os.path.join(HERE, "cornel_textured_ref_48.npz"), **out)
    print("wrote cornel_textured_ref_48.npz", len(out), "entries")


def main():
    if "--lucy" in sys.argv:
        return lucy()
    if "--textured" in sys.argv:
        return textured()
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
