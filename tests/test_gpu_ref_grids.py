"""Counter / heat-map parity on the reference's own Subgrid hierarchy.  The reference derives its grids
from its two-level BVH (scene -> objects -> triangles, 4 BVH levels per grid level, lightFieldGrid.cpp:189-322); with them
uploaded (ptb_upload_subgrid) mode TWO_PLANE walks them like intersectGrid (lightFieldGrid.cpp:71-147), so
getIntersectionCount / getStructureIntersectionCount (tracer.cpp:60-70) and the debug view (tracer.cpp:136-142) are the CPU
tracer's numbers — integers, exactly — for the structure the thesis is about.  The grids come from the oracle, which is pinned
to the compiled reference counters included (tests/test_oracle_pins.py); the 64x64 case below is compared with the golden the
unmodified reference itself wrote (tests/golden/cornel_ref_64.npz: cnt2_nee{0,1})."""
import os

import numpy as np
import pytest

from conftest import GOLDEN
from oracle_lib import MODE_TWO_PLANE, Oracle

pytestmark = pytest.mark.gpu
GS, OS, M = 10, 5, 40  # the parameters of the committed golden (make_golden.py: ptref's defaults)


def _bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


@pytest.fixture(scope="module")
def pair(ptb, cornel_ptbs):
    o = Oracle(cornel_ptbs)
    o.build(grid=True, gs=GS, os_=OS, m=M)
    grids, ls, le, idx = o.export_grids()
    pt = ptb.PathTracer(0)
    pt.load_ptbs(cornel_ptbs)
    pt.build_bvh()
    pt.upload_subgrid(grids.view(ptb.REF_GRID), ls, le, idx)
    yield o, pt, (grids, ls, le, idx)
    pt.close()
    o.close()


@pytest.mark.parametrize("nee", [0, 1])
def test_counters_equal_the_unmodified_references_golden(pair, nee):
    _, pt, _ = pair
    gold = np.load(os.path.join(GOLDEN, "cornel_ref_64.npz"))
    pt.set_params(width=64, height=64, mode=MODE_TWO_PLANE, nee=nee, max_depth=0, count_tests=1)
    pt.reset()
    pt.render(2)
    c = pt.counters()
    img = pt.read_image()
    want = [int(v) for v in gold[f"cnt2_nee{nee}"]]
    print(f"2Plane nee {nee}: triangle tests {c['tri_tests']} (reference {want[0]}), AS tests {c['as_tests']} (reference {want[1]})")
    assert [c["tri_tests"], c["as_tests"]] == want
    assert (_bits(img) == _bits(gold[f"img2_nee{nee}"])).all()


@pytest.mark.parametrize("nee,max_depth", [(0, 0), (1, 0), (0, 5), (1, 3)])
def test_counters_equal_the_oracles(pair, nee, max_depth):
    o, pt, _ = pair
    W, H, spp = 128, 96, 2
    o.configure(W, H, MODE_TWO_PLANE, nee=nee, max_depth=max_depth)
    rimg, rs, rrng, st = o.render(spp)
    pt.set_params(width=W, height=H, mode=MODE_TWO_PLANE, nee=nee, max_depth=max_depth, count_tests=1)
    pt.reset()
    pt.render(spp)
    c = pt.counters()
    s, rng = pt.read_state()
    img = pt.read_image()
    assert c["tri_tests"] == st["tri_tests"] and c["as_tests"] == st["as_tests"]
    assert c["paths"] == st["paths"] == W * H * spp
    if nee == 0:
        assert c["ref_rays"] == st["rays_traversed"]
    assert (s == rs).all() and (rng == rrng).all()
    assert (_bits(img) == _bits(rimg)).all()


def test_heat_maps_equal_per_pixel_counts(pair):
    """tracer.cpp:136-142 over the reference's grids: the pixel is count / scale of its first ray (red above the scale)."""
    o, pt, _ = pair
    W, H, scale = 96, 64, 40
    o.configure(W, H, MODE_TWO_PLANE, nee=0, max_depth=0)
    n_as, n_tri = o.primary_counts()
    for view, cnt in ((1, n_as), (2, n_tri)):
        pt.set_params(width=W, height=H, mode=MODE_TWO_PLANE, nee=0, max_depth=0, debug_view=view, debug_scale=scale, count_tests=1)
        pt.reset()
        pt.render(1)
        img = pt.read_image()[..., :3].reshape(-1, 3)
        it = cnt.astype(np.float32)
        want = np.where((it > scale)[:, None], np.array([1.0, 0.0, 0.0], np.float32)[None, :], (it / np.float32(scale))[:, None].repeat(3, 1))
        assert (_bits(img) == _bits(want.astype(np.float32))).all()
    pt.set_params(width=W, height=H, mode=MODE_TWO_PLANE, nee=0, max_depth=0, debug_view=0, count_tests=0)


def test_own_grids_give_the_same_frame_and_other_counts(ptb, cornel_ptbs, pair):
    _, pt, _ = pair
    W, H = 128, 96
    pt.set_params(width=W, height=H, mode=MODE_TWO_PLANE, nee=1, max_depth=0, count_tests=1)
    pt.reset()
    pt.render(2)
    ref_c, ref_img = pt.counters(), pt.read_image().copy()
    ref_bytes = int(pt.L.ptb_ref_subgrid_bytes(pt.h))
    assert ref_bytes > 0
    own = ptb.PathTracer(0)
    own.load_ptbs(cornel_ptbs)
    own.build_bvh()
    own.build_subgrid(GS, OS, M)
    assert int(own.L.ptb_ref_subgrid_bytes(own.h)) == 0
    own.set_params(width=W, height=H, mode=MODE_TWO_PLANE, nee=1, max_depth=0, count_tests=1)
    own.reset()
    own.render(2)
    c, img = own.counters(), own.read_image()
    own.close()
    differing = int((_bits(img) != _bits(ref_img)).any(axis=-1).sum())
    print(f"own grids vs reference grids: {differing} of {W * H} pixels differ; AS tests {c['as_tests']} vs {ref_c['as_tests']}")
    assert differing <= W * H // 100  # rays that leave a grid's planes find no cell: which ones depends on the hierarchy
    assert c["as_tests"] != ref_c["as_tests"]


def test_bad_grids_are_rejected(ptb, cornel_ptbs, pair):
    _, _, (grids, ls, le, idx) = pair
    g = grids.view(ptb.REF_GRID)
    pt = ptb.PathTracer(0)
    pt.load_ptbs(cornel_ptbs)
    with pytest.raises(ptb.PtbError, match="ptb_build_bvh"):
        pt.upload_subgrid(g, ls, le, idx)
    pt.build_bvh()
    inner = int(np.flatnonzero(g["has_tris"] == 0)[0])
    base = int(g["n_indices"][:inner].sum())
    bad = idx.copy()
    bad[base] = inner  # a grid that lists itself as a child: a cycle
    with pytest.raises(ptb.PtbError, match="not after its parent"):
        pt.upload_subgrid(g, ls, le, bad)
    bad = idx.copy()
    leaf = int(np.flatnonzero(g["has_tris"] == 1)[0])
    bad[int(g["n_indices"][:leaf].sum())] = 1 << 30
    with pytest.raises(ptb.PtbError, match="triangle id out of range"):
        pt.upload_subgrid(g, ls, le, bad)
    bad_le = le.copy()
    bad_le[0] = 0xFFFFFFF0
    with pytest.raises(ptb.PtbError, match="LUT cell outside"):
        pt.upload_subgrid(g, ls, bad_le, idx)
    g2 = g.copy()
    g2["max"][5] += 1.0
    with pytest.raises(ptb.PtbError, match=r"max\[axis\]"):
        pt.upload_subgrid(g2, ls, le, idx)
    with pytest.raises(ptb.PtbError, match="pools"):
        pt.upload_subgrid(g, ls[:-1], le[:-1], idx)
    pt.upload_subgrid(g, ls, le, idx)  # and the intact set still loads
    pt.set_params(width=32, height=32, mode=MODE_TWO_PLANE, nee=0, max_depth=3)
    pt.reset()
    pt.render(1)
    pt.close()
