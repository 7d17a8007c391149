"""The C ABI must answer bad input with a status and a message, never with a C++ exception across extern "C", an
out-of-bounds read on the device or state left over from the previous scene."""
import os
import struct

import numpy as np
import pytest

from oracle_lib import MODE_BVH

pytestmark = pytest.mark.gpu


def test_corrupt_ptbs_header_is_an_io_error(ptb, cornel_ptbs, tmp_path):
    raw = bytearray(open(cornel_ptbs, "rb").read())
    pt = ptb.PathTracer(0)
    # n_tris is the u32 after magic + version (include/ptbs_format.h): claim more than 2^27 triangles, then 2^26 (more than the file holds)
    for claim in (0xFFFFFFF0, 1 << 26):
        bad = bytearray(raw)
        struct.pack_into("<I", bad, 8, claim)
        p = str(tmp_path / f"bad{claim}.ptbs")
        open(p, "wb").write(bad)
        with pytest.raises(ptb.PtbError, match="corrupt .ptbs header"):
            pt.load_ptbs(p)
    p = str(tmp_path / "short.ptbs")
    open(p, "wb").write(raw[: len(raw) // 2])
    with pytest.raises(ptb.PtbError, match="corrupt|truncated"):
        pt.load_ptbs(p)
    pt.load_ptbs(cornel_ptbs)  # the context is still usable
    pt.build_bvh()
    pt.set_params(width=32, height=32, mode=MODE_BVH, nee=0, max_depth=3)
    pt.reset()
    pt.render(1)
    pt.close()


def test_material_index_out_of_range_is_rejected(ptb, cornel_ptbs):
    sc = ptb.scenes
    scene = sc.read_ptbs(cornel_ptbs)
    mat = scene["mat"].copy()
    mat[5] = len(scene["materials"])  # one past the table
    pt = ptb.PathTracer(0)
    sc.upload_scene(pt, dict(scene, mat=mat))
    pt.build_bvh()
    pt.set_params(width=32, height=32, mode=MODE_BVH, nee=0, max_depth=3)
    with pytest.raises(ptb.PtbError, match="material index"):
        pt.render(1)
    with pytest.raises(ptb.PtbError, match="material index"):
        pt.primary_hits()
    pt.upload_triangles(scene["pos"], scene["nrm"], scene["uv"], scene["mat"])  # repaired
    pt.build_bvh()
    pt.reset()
    pt.render(1)
    pt.close()


def test_untextured_scene_after_a_textured_one_renders_untextured(ptb, cornel_ptbs, tmp_path):
    """Same material count, no textures: the bindings of the previous scene must be gone (they used to survive)."""
    sc = ptb.scenes
    room = sc.read_ptbs(cornel_ptbs)
    tex = sc.textured_cornell(room, light_texture=False)
    plain = dict(tex, textures=[], mat_tex=None)
    p_tex, p_plain = str(tmp_path / "tex.ptbs"), str(tmp_path / "plain.ptbs")
    sc.write_ptbs(p_tex, tex)
    sc.write_ptbs(p_plain, plain)

    def frame(pt):
        pt.build_bvh()
        pt.set_params(width=64, height=48, mode=MODE_BVH, nee=1, max_depth=4)
        pt.reset()
        pt.render(2)
        return pt.read_image().copy()

    fresh = ptb.PathTracer(0)
    fresh.load_ptbs(p_plain)
    want = frame(fresh)
    fresh.close()
    pt = ptb.PathTracer(0)
    pt.load_ptbs(p_tex)
    textured = frame(pt)
    pt.load_ptbs(p_plain)
    got = frame(pt)
    pt.close()
    assert not (textured.view(np.uint32) == want.view(np.uint32)).all(), "the textured scene should differ"
    assert (got.view(np.uint32) == want.view(np.uint32)).all()
