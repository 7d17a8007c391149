"""ptb_load_scene_file on the GPU: scene text (.scene + OBJ + MTL + a PNG texture, written by the test) loaded directly must
render exactly what the same text converted to .ptbs (ptb_convert_scene_file, pinned against the reference's loader in
tests/test_scene_loader.py) and loaded with ptb_load_ptbs renders, and what the oracle renders from that container."""
import os

import numpy as np
import pytest

from oracle_lib import MODE_BVH, Oracle
from test_scene_loader import _png_bytes

pytestmark = pytest.mark.gpu


def _bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def _tree(d):
    os.makedirs(os.path.join(d, "objs"))
    rng = np.random.default_rng(2)
    with open(os.path.join(d, "objs", "checker.png"), "wb") as f:
        f.write(_png_bytes(rng.integers(0, 256, (8, 8, 4)), 6, 8, [0, 1, 2, 3, 4]))
    with open(os.path.join(d, "objs", "m.mtl"), "w") as f:
        f.write("newmtl wall\nKd 0.7 0.7 0.7\nmap_Kd checker.png\nnewmtl lamp\nKd 1 1 1\nKe 12 12 10\nnewmtl floor\nKd 0.2 0.6 0.3\n")
    with open(os.path.join(d, "objs", "room.obj"), "w") as f:
        f.write("mtllib m.mtl\n"
                "v -1 -1 1\nv 1 -1 1\nv 1 1 1\nv -1 1 1\nv -1 -1 -1\nv 1 -1 -1\nv -0.3 0.95 0.2\nv 0.3 0.95 0.2\nv 0.3 0.95 0.8\nv -0.3 0.95 0.8\n"
                "vn 0 0 -1\nvn 0 1 0\nvn 0 -1 0\nvt 0 0\nvt 1 0\nvt 1 1\nvt 0 1\n"
                "o back\nusemtl wall\nf 1/1/1 2/2/1 3/3/1 4/4/1\n"
                "o ground\nusemtl floor\nf 5/1/2 6/2/2 2/3/2 1/4/2\n"
                "o light\nusemtl lamp\nf 7/1/3 8/2/3 9/3/3 10/4/3\n")
    with open(os.path.join(d, "s.scene"), "w") as f:
        f.write("[Models]\nfile: objs/room.obj\npos: 0,0,0\nscale: 100,100,100\n\n[Camera]\npos: 0,0,-300\nforward: 0,0,1\nfocus: 1\n")
    return os.path.join(d, "s.scene")


def test_scene_text_renders_like_its_ptbs(ptb, tmp_path):
    d = str(tmp_path)
    scene, ptbs = _tree(d), os.path.join(d, "s.ptbs")
    ptb.convert_scene_file(scene, ptbs, base_dir=d)
    W, H, spp = 96, 72, 3
    frames = []
    for direct in (True, False):
        pt = ptb.PathTracer(0)
        if direct:
            pt.load_scene_file(scene, d)
        else:
            pt.load_ptbs(ptbs)
        pt.build_bvh()
        pt.set_params(width=W, height=H, mode=MODE_BVH, nee=1, max_depth=4)
        pt.reset()
        pt.render(spp)
        frames.append((pt.read_image(), pt.read_state()))
        pt.close()
    (a, (sa, ra)), (b, (sb, rb)) = frames
    assert (_bits(a) == _bits(b)).all() and (sa == sb).all() and (ra == rb).all()
    o = Oracle(ptbs)
    o.build()
    o.configure(W, H, MODE_BVH, nee=1, max_depth=4)
    rimg, rs, rrng, _ = o.render(spp)
    o.close()
    assert float(rimg[..., :3].max()) > 0  # the lamp lights the textured wall
    assert (sa == rs).all() and int((ra != rrng).sum()) <= W * H // 2000
    assert int((_bits(a) != _bits(rimg)).any(axis=-1).sum()) <= W * H // 2000
