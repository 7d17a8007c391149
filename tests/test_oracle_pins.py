"""Pins the CPU oracle (oracle/ptoracle.cpp -> libptoracle.so) to the reference.

1. Against the committed golden vectors (tests/golden/cornel_ref_64.npz), which were produced by
   the UNMODIFIED reference compiled in the build container (tests/golden/make_golden.py): primary
   hit distance / triangle id / material, images after 2 spp with and without NEE, the reference's
   two test counters, and depth-capped images — bit-exact, all four intersect modes.
2. Where oracle/_ref/ptref exists (build container; it also travels to the GPU box), live against
   it at another resolution and sample count, so the pin is not limited to the fixture size.
3. The reference's own cross-mode consistency: ALL == BVH == TWO_PLANE on tmax.
"""
import os
import tempfile

import numpy as np
import pytest

from oracle_lib import MODE_ALL, MODE_BVH, MODE_SUB_BEAMS, MODE_TWO_PLANE, Oracle, have_ref, ptref_primary_hits, ptref_render, run_ptref

MODES = [MODE_ALL, MODE_BVH, MODE_TWO_PLANE, MODE_SUB_BEAMS]
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "cornel_ref_64.npz")


@pytest.fixture(scope="module")
def gold():
    return np.load(GOLD)


@pytest.fixture(scope="module")
def oracle(cornel_ptbs):
    o = Oracle(cornel_ptbs)
    o.build(grid=True, beams=True)  # reference defaults: gs 10, os 5, m 40 (main.cpp:42-44)
    yield o
    o.close()


def _bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


@pytest.mark.parametrize("mode", MODES)
def test_primary_hits_match_reference_golden(oracle, gold, mode):
    oracle.configure(64, 64, mode)
    t, tri, mat, rays = oracle.primary_hits(want_rays=True)
    assert (_bits(t) == _bits(gold[f"t{mode}"])).all()
    assert (tri == gold[f"tri{mode}"]).all()
    assert (mat == gold[f"mat{mode}"]).all()
    assert (_bits(rays) == _bits(gold["rays"])).all()


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("nee", [0, 1])
def test_images_and_counters_match_reference_golden(oracle, gold, mode, nee):
    oracle.configure(64, 64, mode, nee=nee, max_depth=0)
    img, samples, rng, st = oracle.render(2)
    assert (samples == 2).all()
    assert (_bits(img) == _bits(gold[f"img{mode}_nee{nee}"])).all()
    cnt = gold[f"cnt{mode}_nee{nee}"]
    assert st["tri_tests"] == int(cnt[0]) and st["as_tests"] == int(cnt[1])


@pytest.mark.parametrize("mode", MODES)
def test_depth_capped_image_matches_reference_golden(oracle, gold, mode):
    oracle.configure(64, 64, mode, nee=1, max_depth=2)
    img, _, _, _ = oracle.render(3)
    assert (_bits(img) == _bits(gold[f"imgd{mode}"])).all()


def test_cross_mode_tmax_identical(oracle):
    ts = []
    for mode in (MODE_ALL, MODE_BVH, MODE_TWO_PLANE):
        oracle.configure(96, 96, mode)
        ts.append(oracle.primary_hits()[0])
    assert (_bits(ts[0]) == _bits(ts[1])).all() and (_bits(ts[0]) == _bits(ts[2])).all()


def test_progressive_continuation(oracle):
    oracle.configure(48, 48, MODE_BVH, nee=1)
    a, _, _, _ = oracle.render(3)
    b, s, r, _ = oracle.render(1)
    b2, _, _, _ = oracle.render(2, state=(b.copy(), s, r))
    assert (_bits(a) == _bits(b2)).all()


@pytest.mark.skipif(not have_ref(), reason="oracle/_ref/ptref not built (needs the reference sources)")
@pytest.mark.parametrize("mode", [MODE_BVH, MODE_TWO_PLANE, MODE_SUB_BEAMS])
def test_live_against_compiled_reference(oracle, cornel_ptbs, mode):
    W, H = 96, 80
    oracle.configure(W, H, mode, nee=1, max_depth=5)
    t, tri, mat = oracle.primary_hits()
    rt, rtri, rmat, _, _ = ptref_primary_hits(cornel_ptbs, W, H, mode)
    assert (_bits(t) == _bits(rt)).all() and (tri == rtri).all() and (mat == rmat).all()
    img, _, _, st = oracle.render(3)
    rimg, rec, _ = ptref_render(cornel_ptbs, W, H, mode, 3, 5, 1)
    assert (_bits(img) == _bits(rimg)).all()
    assert st["tri_tests"] == rec["tri_tests"] and st["as_tests"] == rec["as_tests"]


# ---- texture branches (shader.cpp:52-66 albedo texture + alpha, :251-258 normal map; texture.cpp:22-32) ------------
TEX_GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "cornel_textured_ref_48.npz")


@pytest.fixture(scope="module")
def textured_scenes(cornel_ptbs, tmp_path_factory):
    """scenes.textured_cornell with and without a texture on the emissive cube, as .ptbs v2 files; the generator must
    reproduce the bytes the golden images were rendered from."""
    import hashlib
    import ptb_b200
    sc = ptb_b200.scenes
    gold = np.load(TEX_GOLD)
    d = tmp_path_factory.mktemp("textured")
    paths = {}
    for lt in (0, 1):
        scene = sc.textured_cornell(sc.read_ptbs(cornel_ptbs), light_texture=bool(lt))
        h = hashlib.sha256(scene["uv"].tobytes())
        for t in scene["textures"]:
            h.update(np.ascontiguousarray(t, np.float32).tobytes())
        assert h.hexdigest() == str(gold[f"scene_sha256_l{lt}"]), "textured scene generator is not bit-reproducible here"
        paths[lt] = str(d / f"tex{lt}.ptbs")
        sc.write_ptbs(paths[lt], scene)
        back = sc.read_ptbs(paths[lt])
        assert all((a == b).all() for a, b in zip(back["textures"], scene["textures"]))
        assert [(m.get("tex", -1), m.get("nrm", -1)) for m in back["materials"]] == [(m.get("tex", -1), m.get("nrm", -1)) for m in scene["materials"]]
    return paths, gold


@pytest.mark.parametrize("lt", [0, 1])
@pytest.mark.parametrize("mode", [MODE_BVH, MODE_TWO_PLANE])
def test_textured_images_match_reference_golden(textured_scenes, mode, lt):
    paths, gold = textured_scenes
    o = Oracle(paths[lt])
    o.build(grid=True)
    for nee, depth in ((0, 0), (1, 0), (1, 4)):
        o.configure(64, 48, mode, nee=nee, max_depth=depth)
        img, _, _, _ = o.render(3)
        assert (_bits(img) == _bits(gold[f"img_m{mode}_nee{nee}_d{depth}_l{lt}"])).all(), f"mode {mode} nee {nee} depth {depth} light texture {lt}"
    o.close()


@pytest.mark.skipif(not have_ref(), reason="oracle/_ref/ptref not built (needs the reference sources)")
def test_textured_live_against_compiled_reference(textured_scenes):
    paths, _ = textured_scenes
    W, H = 80, 56
    o = Oracle(paths[1])
    o.build()
    o.configure(W, H, MODE_BVH, nee=1, max_depth=6)
    img, _, _, st = o.render(2)
    o.close()
    rimg, rec, _ = ptref_render(paths[1], W, H, MODE_BVH, 2, 6, 1)
    assert (_bits(img) == _bits(rimg)).all()
    assert st["tri_tests"] == rec["tri_tests"] and st["as_tests"] == rec["as_tests"]


# ---- a second scene: two models, one of them a flat Blender export (zero-thickness boxes in every structure) -----------------
TWO_MODELS = """[Models]
file: objects/cornel.obj
pos:  0,-250,0
scale: 300,300,300
file: objects/test.obj
pos: 40,-60,-120
scale: 18,14,300

[Materials]
name: Material.004
color: 1.0,1.0,0.8
emission: 0
roughness: 0.1
weights: 0,0,1

[Camera]
pos: 0,0,-500
forward: 0,0.0,1
focus: 1
"""


@pytest.mark.skipif(not (have_ref() and os.path.isfile("/root/reference/showcase/objects/test.obj")), reason="needs ptref and the reference's showcase/ assets")
@pytest.mark.parametrize("mode", [MODE_BVH, MODE_TWO_PLANE])
def test_live_two_model_scene_with_a_flat_mesh(tmp_path, mode):
    """Cornell plus showcase/objects/test.obj (1,932 coplanar triangles, 23 % of the frame) loaded from scene text: hits, frame
    and both counters of the restatement equal the compiled reference's.  (Subbeams are left out: the reference's beam
    recursion has no depth guard and does not terminate on this mesh)"""
    import ptb_b200
    scene, ptbs = str(tmp_path / "two.scene"), str(tmp_path / "two.ptbs")
    with open(scene, "w") as f:
        f.write(TWO_MODELS)
    ptb_b200.convert_scene_file(scene, ptbs, base_dir="/root/reference/showcase")
    W, H = 96, 80
    o = Oracle(ptbs)
    o.build(grid=(mode == MODE_TWO_PLANE))
    o.configure(W, H, mode, nee=1, max_depth=5)
    t, tri, mat = o.primary_hits()
    img, _, _, st = o.render(2)
    o.close()
    rt, rtri, rmat, _, _ = ptref_primary_hits(ptbs, W, H, mode)
    rimg, rec, _ = ptref_render(ptbs, W, H, mode, 2, 5, 1)
    assert (_bits(t) == _bits(rt)).all() and (tri == rtri).all() and (mat == rmat).all()
    assert 0.1 < float((mat == mat.max()).mean()) < 0.5  # the flat mesh is in view
    assert (_bits(img) == _bits(rimg)).all()
    assert st["tri_tests"] == rec["tri_tests"] and st["as_tests"] == rec["as_tests"]


@pytest.mark.skipif(not have_ref(), reason="oracle/_ref/ptref not built (needs the reference sources)")
@pytest.mark.parametrize("mode", [MODE_BVH, MODE_TWO_PLANE])
def test_live_from_scene_text_end_to_end(tmp_path, mode):
    """The whole chain against the unmodified reference started from TEXT: the reference reads .scene / OBJ / MTL / PNG with
    its own loader (tinyobjloader, stb_image) and renders; the restatement renders the .ptbs this repo's loader wrote from the
    same text (tests/test_gpu_scene_file.py's room: a PNG-textured wall, a lamp, quads).  Frames and counters must agree."""
    import ptb_b200
    from test_gpu_scene_file import _tree
    d = str(tmp_path)
    scene, ptbs = _tree(d), os.path.join(d, "s.ptbs")
    ptb_b200.convert_scene_file(scene, ptbs, base_dir=d)
    W, H, spp = 64, 48, 2
    o = Oracle(ptbs)
    o.build(grid=(mode == MODE_TWO_PLANE))
    o.configure(W, H, mode, nee=1, max_depth=4)
    img, _, _, st = o.render(spp)
    o.close()
    with tempfile.TemporaryDirectory() as td:
        out = os.path.join(td, "img.bin")
        recs = run_ptref(["--scene-file", scene, "--base", d, "--mode", mode, "--width", W, "--height", H, "--spp", spp, "--max-depth", 4,
                          "--nee", 1, "render", out])
        rimg = np.fromfile(out, np.float32).reshape(H, W, 4)
    rec = [r for r in recs if r["cmd"] == "render"][0]
    assert float(rimg[..., :3].max()) > 0
    assert (_bits(img) == _bits(rimg)).all()
    assert st["tri_tests"] == rec["tri_tests"] and st["as_tests"] == rec["as_tests"]
