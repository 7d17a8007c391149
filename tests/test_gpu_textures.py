"""Texture branches of the shader on the GPU (shade_hit<.., TEX>, k_mega<.., TEX>): getColorOfMaterial
(shader.cpp:52-66: nearest texel, alpha draw, see-through step), the normal map (shader.cpp:251-258), the wrap rule
of getTextureAtUV (texture.cpp:22-32) and a textured light under NEE (shader.cpp:104-109) — against the reference's own
golden images (tests/golden/cornel_textured_ref_48.npz, rendered by the unmodified reference) and the oracle.
Same bar as everywhere: frames, RNG streams and sample counts bit-equal (<= W*H/2000 pixels may differ at all)."""
import os

import numpy as np
import pytest

from oracle_lib import MODE_BVH, MODE_SUB_BEAMS, MODE_TWO_PLANE, Oracle

pytestmark = pytest.mark.gpu
TEX_GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "cornel_textured_ref_48.npz")


def _bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def _ndiff(a, b):
    return int((_bits(a) != _bits(b)).any(axis=-1).sum())


@pytest.fixture(scope="module")
def scenes(ptb, cornel_ptbs, tmp_path_factory):
    sc = ptb.scenes
    d = tmp_path_factory.mktemp("textured")
    out = {}
    for lt in (0, 1):
        scene = sc.textured_cornell(sc.read_ptbs(cornel_ptbs), light_texture=bool(lt))
        path = str(d / f"tex{lt}.ptbs")
        sc.write_ptbs(path, scene)
        out[lt] = (scene, path)
    return out


@pytest.mark.parametrize("lt", [0, 1])
def test_textured_frames_equal_the_references_golden(ptb, scenes, lt):
    gold = np.load(TEX_GOLD)
    scene, path = scenes[lt]
    W, H = 64, 48
    pt = ptb.PathTracer(0)
    pt.load_ptbs(path)  # the v2 container: textures + bindings arrive through ptb_load_ptbs
    pt.build_bvh()
    pt.build_subgrid(10, 5, 40)
    for mode in (MODE_BVH, MODE_TWO_PLANE):
        for nee, depth in ((0, 0), (1, 0), (1, 4)):
            pt.set_params(width=W, height=H, mode=mode, nee=nee, max_depth=depth)
            pt.reset()
            pt.render(3)
            img = pt.read_image()
            n = _ndiff(img, gold[f"img_m{mode}_nee{nee}_d{depth}_l{lt}"])
            assert n <= max(1, W * H // 2000), f"mode {mode} nee {nee} depth {depth} light texture {lt}: {n} pixels differ from the reference"
    pt.close()


def test_textured_frame_vs_oracle_all_structures(ptb, scenes):
    scene, path = scenes[1]
    W, H, spp = 160, 120, 2
    o = Oracle(path)
    o.build(grid=True, beams=True)
    pt = ptb.PathTracer(0)
    ptb.scenes.upload_scene(pt, scene)  # textures through ptb_upload_textures
    pt.build_bvh(gpu=True)
    pt.build_subgrid(10, 5, 40, gpu=True)
    pt.build_subbeams(10, 40)
    frames = {}
    for mode in (MODE_BVH, MODE_TWO_PLANE, MODE_SUB_BEAMS):
        o.configure(W, H, mode, nee=1, max_depth=7)
        rimg, rs, rrng, st = o.render(spp)
        pt.set_params(width=W, height=H, mode=mode, nee=1, max_depth=7)
        pt.reset()
        pt.render(spp)
        s, rng = pt.read_state()
        img = pt.read_image()
        assert (s == rs).all() and int((rng != rrng).sum()) <= W * H // 2000, f"mode {mode}: RNG streams diverged"
        assert _ndiff(img, rimg) <= W * H // 2000, f"mode {mode}: {_ndiff(img, rimg)} pixels differ from the oracle"
        assert pt.counters()["paths"] == st["paths"]
        frames[mode] = img
    # the UI flow: materials re-uploaded before a reset keep their texture bindings
    sc = ptb.scenes
    pt.upload_materials(sc.pack_materials(scene["materials"]), len(scene["materials"]))
    pt.set_params(mode=MODE_BVH)
    pt.reset()
    pt.render(spp)
    assert (_bits(pt.read_image()) == _bits(frames[MODE_BVH])).all()
    # and without the textures the same scene renders differently (the branches are live)
    pt.upload_textures([], np.full((len(scene["materials"]), 2), -1, np.int32))
    pt.reset()
    pt.render(spp)
    assert _ndiff(pt.read_image(), frames[MODE_BVH]) > W * H // 20
    pt.close()
    o.close()


def test_untextured_scene_is_unaffected_and_errors_are_loud(ptb, cornel_ptbs):
    pt = ptb.PathTracer(0)
    pt.load_ptbs(cornel_ptbs)
    pt.build_bvh()
    pt.set_params(width=64, height=64, mode=MODE_BVH, nee=1, max_depth=4)
    pt.reset()
    pt.render(2)
    a = pt.read_image()
    tex = [np.ones((2, 2, 4), np.float32)]
    with pytest.raises(ptb.PtbError):  # binding out of range
        pt.upload_textures(tex, np.full((8, 2), 3, np.int32))
    with pytest.raises(ptb.PtbError):  # one binding per material
        pt.upload_textures(tex, np.full((3, 2), -1, np.int32))
    pt.upload_textures(tex, np.full((8, 2), -1, np.int32))  # textures present, nothing bound: the untextured kernel
    pt.reset()
    pt.render(2)
    assert (_bits(pt.read_image()) == _bits(a)).all()
    pt.close()
