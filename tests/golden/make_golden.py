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

  tests/golden/configs_ref.npz         (--configs [--lucy-subgrid]) the BASELINE.json configurations at their own sizes
        c0_tri, c0_t_sha256           config 0: Cornell 512x512, BVH: primary hit ids and the digest of the hit distances
        c0_img_sha256, c0_thumb, c0_counts    ... the fp32 frame after 1 spp, max_depth 2, NEE off: digest of the bit
                                      pattern, 8x8 block means (for a readable diff), [find_calls, rays_traversed,
                                      rays_unique, tri_tests, as_tests] of the unmodified trace(true)
        c1_*                          config 1 at the first 4 of its 64 spp: Cornell 1024x1024, max_depth 5
        lucy_bvh_*                    Lucy stand-in 1920x1080, 2 spp, max_depth 9, BVH (scene digest: lucy_scene_sha256)
        lucy_grid_*                   (--lucy-subgrid, 14 min / 11 GB) the same through the reference's Subgrids gs 8 os 8 m 50;
                                      lucy_grid_img_f16 = the frame in half precision (the GPU builds its grids from another
                                      hierarchy, so this frame is compared with a stated tolerance, not by digest)
  tests/golden/cornel_rmse_ref.npz     (--rmse) config 1's converged-image bar, see rmse_golden()

    python tests/golden/make_golden.py [--lucy [--lucy-two-plane]] [--textured] [--configs [--lucy-subgrid]] [--rmse]
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


def thumb(img, k=8):
    h, w = img.shape[0] // k * k, img.shape[1] // k * k
    return img[:h, :w, :3].astype(np.float64).reshape(h // k, k, w // k, k, 3).mean(axis=(1, 3)).astype(np.float32)


def counts(rec):
    return np.array([rec[k] for k in ("find_calls", "rays_traversed", "rays_unique", "tri_tests", "as_tests")], np.uint64)


def configs():
    import hashlib
    import tempfile
    sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
    import ptb_b200
    sc = ptb_b200.scenes
    path = os.path.join(HERE, "configs_ref.npz")
    out = dict(np.load(path)) if os.path.isfile(path) else {}
    cornel = os.path.join(HERE, "cornel.ptbs")
    sha = lambda a: hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()

    def frame(prefix, ptbs, W, H, spp, depth, mode=1, extra=()):
        img, rec, _ = ptref_render(ptbs, W, H, mode, spp, depth, 0, extra=extra)
        out[prefix + "_img_sha256"] = sha(img)
        out[prefix + "_thumb"] = thumb(img)
        out[prefix + "_counts"] = counts(rec)
        out[prefix + "_trace_ms"] = rec["trace_ms"]
        print(prefix, out[prefix + "_img_sha256"], out[prefix + "_counts"], rec["trace_ms"], "ms")
        return img

    if "--lucy-subgrid" not in sys.argv:
        t, tri, mat, rays, _ = ptref_primary_hits(cornel, 512, 512, 1)
        out["c0_tri"], out["c0_t_sha256"] = tri, sha(t)
        frame("c0", cornel, 512, 512, 1, 2)
        frame("c1", cornel, 1024, 1024, 4, 5)
    scene = sc.lucy_standin(sc.read_ptbs(cornel))
    out["lucy_scene_sha256"] = hashlib.sha256(scene["pos"].tobytes()).hexdigest()
    with tempfile.TemporaryDirectory() as d:
        ptbs = os.path.join(d, "lucy.ptbs")
        sc.write_ptbs(ptbs, scene)
        if "--lucy-subgrid" in sys.argv:
            img = frame("lucy_grid", ptbs, 1920, 1080, 2, 9, mode=2, extra=["--gs", 8, "--os", 8, "--m", 50])
            out["lucy_grid_img_f16"] = np.clip(img[..., :3], 0, 65000).astype(np.float16)
        else:
            frame("lucy_bvh", ptbs, 1920, 1080, 2, 9)
    np.savez_compressed(path, **out)
    print("wrote", path, sorted(out))


def rmse_golden():
    """config 1's converged-image bar: the reference's own 4096-spp frame (stored in half precision, radiance
    clamped to 64) and the RMSE of its 64-spp frame against it.  512x512 instead of 1024x1024: 4096 spp of the reference is
    18 CPU-minutes at this size on 8 cores, 70 at the full size; the bar is per pixel."""
    path = os.path.join(HERE, "cornel_rmse_ref.npz")
    cornel = os.path.join(HERE, "cornel.ptbs")
    W = H = 512
    img64, _, _ = ptref_render(cornel, W, H, 1, 64, 5, 0)
    ref, rec, _ = ptref_render(cornel, W, H, 1, 4096, 5, 0)
    a, b = np.minimum(img64[..., :3], 64.0).astype(np.float64), np.minimum(ref[..., :3], 64.0).astype(np.float64)
    out = {"ref4096_f16": np.minimum(ref[..., :3], 64.0).astype(np.float16), "rmse_cpu64": float(np.sqrt(((a - b) ** 2).mean())),
           "ref_trace_ms": rec["trace_ms"], "size": np.array([W, H]), "spp": np.array([64, 4096]), "max_depth": 5}
    np.savez_compressed(path, **out)
    print("wrote", path, out["rmse_cpu64"], rec["trace_ms"])


def main():
    if "--rmse" in sys.argv:
        return rmse_golden()
    if "--configs" in sys.argv:
        return configs()
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
