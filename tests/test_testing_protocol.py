"""The reference's `--testing` protocol, headless (testing_protocol.py; main.cpp:99-141, window.cpp:540-546): the five stdout
lines per structure in the form showcase/test.py parses, the PNG names, and — with the reference's own tree AND its own grids
uploaded — the counters of the CPU tracer in both blocks, and the PNGs byte for byte what the reference's window would show."""
import os
import re

import numpy as np
import pytest

from conftest import apply_display
from oracle_lib import MODE_BVH, MODE_TWO_PLANE, Oracle

PATTERN = re.compile(r'(\w+ \w+): (\d{2}:\d{2}:\d{2}|\d+)')  # showcase/test.py:13
KEYS = ["Rendering Time", "Triangle Intersections", "AS Intersections", "Total Intersections", "Memory Consumption"]


def test_png_round_trip_and_names(tmp_path):
    import ptb_b200
    tp = ptb_b200.testing_protocol
    a = np.random.default_rng(5).integers(0, 256, (31, 47, 4)).astype(np.uint8)
    p = str(tmp_path / "a.png")
    tp.write_png(p, a)
    assert (tp.read_png(p) == a).all()
    with pytest.raises(ValueError):
        tp.write_png(p, a[..., :3])
    assert [tp.file_name(1, False, False), tp.file_name(1, True, False), tp.file_name(2, True, True), tp.file_name(3, False, False)] == \
        ["render_BVH.png", "render_heatmap_as_BVH.png", "render_heatmap_tris_Subgrids.png", "render_Subbeams.png"]


@pytest.mark.gpu
def test_protocol_output_counters_and_files(ptb, cornel_ptbs, tmp_path, tonemap_checker):
    W, H, S = 160, 90, 3
    o = Oracle(cornel_ptbs)
    o.build(grid=True, gs=10, os_=5, m=40)
    nodes, idx, roots, root = o.export_bvh()
    grids, lut_s, lut_e, gidx = o.export_grids()
    o.configure(W, H, MODE_TWO_PLANE, nee=0, max_depth=0)
    grid_img, _, _, gst = o.render(S)
    ga_cnt, gt_cnt = o.primary_counts()
    o.configure(W, H, MODE_BVH, nee=0, max_depth=0)
    bvh_img, _, _, st = o.render(S)
    a_cnt, t_cnt = o.primary_counts()
    o.close()
    pt = ptb.PathTracer(0)
    pt.load_ptbs(cornel_ptbs)
    lines = []
    recs = ptb.testing_protocol.run(pt, samples=S, gs=10, os_=5, m=40, width=W, height=H, out_dir=str(tmp_path / "render"),
                                    ref_tree=(nodes.view(ptb.REF_BVH_NODE), idx, roots, root),
                                    ref_grids=(grids.view(ptb.REF_GRID), lut_s, lut_e, gidx), emit=lines.append)
    assert lines[0] == "starting test" and len(lines) == 11
    parsed = [PATTERN.findall(l)[0] for l in lines[1:]]
    assert [k for k, _ in parsed] == KEYS + KEYS
    bvh, grid = recs
    assert (bvh["mode"], grid["mode"]) == (MODE_BVH, MODE_TWO_PLANE)
    # BVH pass over the reference's tree: the CPU tracer's numbers (tests/test_oracle_pins.py pins the oracle's to ptref's)
    assert (bvh["tri_tests"], bvh["as_tests"]) == (st["tri_tests"], st["as_tests"])
    assert int(parsed[3][1]) == st["tri_tests"] + st["as_tests"]
    # Subgrids pass over the reference's grids: the other half of the protocol's numbers, equally the CPU tracer's
    assert (grid["tri_tests"], grid["as_tests"]) == (gst["tri_tests"], gst["as_tests"])
    assert int(parsed[6][1]) == gst["tri_tests"] and int(parsed[7][1]) == gst["as_tests"] and int(parsed[8][1]) == gst["tri_tests"] + gst["as_tests"]
    # getMemory2Plane of those grids: 288-byte Grid objects + the three u32 vectors (lightFieldGrid.cpp:60-68)
    assert int(parsed[9][1]) == len(grids) // 80 * 288 + 4 * (2 * len(lut_s) + len(gidx))
    assert int(parsed[4][1]) == pt.structure_bytes(MODE_BVH)
    names = [os.path.basename(f) for r in recs for f in r["files"]]
    assert names == ["render_BVH.png", "render_heatmap_as_BVH.png", "render_heatmap_tris_BVH.png", "render_Subgrids.png",
                     "render_heatmap_as_Subgrids.png", "render_heatmap_tris_Subgrids.png"]
    tp = ptb.testing_protocol
    imgs = [tp.read_png(f) for r in recs for f in r["files"]]
    assert all(i.shape == (H, W, 4) and (i[..., 3] == 255).all() for i in imgs)
    # the renders are the CPU tracer's frames through the display transform (window.cpp:507-531), byte for byte: the checker is
    # the reference's expression compiled with its flags (tests/native/tonemap_check.cpp)
    for png, ref in ((imgs[0], bvh_img), (imgs[3], grid_img)):
        want = apply_display(tonemap_checker, ref, 1, str(tmp_path))
        assert int((png != want).sum()) == 0
    # BVH heat maps are the reference's: grey = count / 10 below the scale, red above (tracer.cpp:136-142), then
    # the display transform without tone mapping (window.cpp:507-531)
    for img, cnt in ((imgs[1], a_cnt), (imgs[2], t_cnt), (imgs[4], ga_cnt), (imgs[5], gt_cnt)):
        over = (cnt.reshape(H, W) > 10)
        assert (img[over][:, 0] == 255).all() and (img[over][:, 1:3] == 0).all()
        assert (img[~over][:, 0] == img[~over][:, 1]).all() and (img[~over][:, 1] == img[~over][:, 2]).all()
        assert (img[cnt.reshape(H, W) == 0][:, :3] == 0).all()
    pt.close()
