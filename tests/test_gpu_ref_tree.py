"""Counter / heat-map parity on the reference's own structure: with the reference's two-level BVH
uploaded (ptb_upload_ref_bvh) and switched on (ptb_use_ref_bvh), mode BVH walks it statement by statement like
findBVHIntesection (bvh.cpp:48-99), so getIntersectionCount / getStructureIntersectionCount (tracer.cpp:60-70) and
the debug view (tracer.cpp:136-142) are the CPU tracer's numbers — exactly, integers — while hits, frames and RNG
streams stay what they were.  The tree comes from the oracle (pinned to the compiled reference, counters included:
tests/test_oracle_pins.py), and in the drop-in flavour from the reference's own accessors (host/tracer_facade.cpp)."""
import numpy as np
import pytest

from oracle_lib import MODE_BVH, Oracle, have_facade, have_ref, ptfacade_render, ptref_render

pytestmark = pytest.mark.gpu


def _bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def _pair(ptb, path, W, H):
    o = Oracle(path)
    o.build()
    nodes, idx, roots, root = o.export_bvh()
    pt = ptb.PathTracer(0)
    pt.load_ptbs(path)
    pt.build_bvh()
    pt.upload_ref_bvh(nodes.view(ptb.REF_BVH_NODE), idx, roots, root)
    pt.use_ref_bvh(True)
    return o, pt


@pytest.mark.parametrize("nee,max_depth", [(0, 0), (1, 0), (0, 5), (1, 3)])
def test_cornell_counters_are_the_references(ptb, cornel_ptbs, nee, max_depth):
    W, H, spp = 128, 96, 2
    o, pt = _pair(ptb, cornel_ptbs, W, H)
    o.configure(W, H, MODE_BVH, nee=nee, max_depth=max_depth)
    rimg, rs, rrng, st = o.render(spp)
    pt.set_params(width=W, height=H, mode=MODE_BVH, nee=nee, max_depth=max_depth, count_tests=1)
    pt.reset()
    pt.render(spp)
    c = pt.counters()
    s, rng = pt.read_state()
    img = pt.read_image()
    # the same frame without the reference tree (own SAH tree): identical pixels, different counts
    pt.use_ref_bvh(False)
    pt.reset()
    pt.render(spp)
    own = pt.counters()
    img2 = pt.read_image()
    pt.close()
    o.close()
    assert c["tri_tests"] == st["tri_tests"], f"triangle tests {c['tri_tests']} != reference {st['tri_tests']}"
    assert c["as_tests"] == st["as_tests"], f"AS tests {c['as_tests']} != reference {st['as_tests']}"
    assert c["paths"] == st["paths"] == W * H * spp
    if nee == 0:  # with NEE the GPU count also holds the shadow rays, which the path-ray statistic of the oracle does not
        assert c["ref_rays"] == st["rays_traversed"]
    assert (s == rs).all() and int((rng != rrng).sum()) <= W * H // 2000
    assert int((_bits(img) != _bits(rimg)).any(axis=-1).sum()) <= W * H // 2000
    assert (_bits(img) == _bits(img2)).all()
    assert own["as_tests"] != c["as_tests"] and own["rays"] == c["rays"]


def test_heat_maps_are_the_references(ptb, cornel_ptbs):
    W, H = 160, 120
    o, pt = _pair(ptb, cornel_ptbs, W, H)
    o.configure(W, H, MODE_BVH)
    a, t = o.primary_counts()
    o.close()
    for view, cnt, scale in ((1, a, 30), (2, t, 6), (1, a, 12)):
        pt.set_params(width=W, height=H, mode=MODE_BVH, nee=0, max_depth=0, debug_view=view, debug_scale=scale, count_tests=1)
        pt.reset()
        pt.render(1)
        img = pt.read_image().reshape(-1, 4)
        it = cnt.astype(np.float32)
        grey = it / np.float32(scale)  # tracer.cpp:138-139
        over = it > np.float32(scale)
        want = np.stack([np.where(over, np.float32(1), grey), np.where(over, np.float32(0), grey), np.where(over, np.float32(0), grey)], axis=-1)
        assert (_bits(img[:, :3]) == _bits(want.astype(np.float32))).all(), f"heat map view {view} scale {scale} differs"
        assert over.any() or scale == 30
    pt.close()


def test_lucy_level6_counters_on_a_deep_object_tree(ptb, cornel_ptbs, tmp_path):
    sc = ptb.scenes
    path = str(tmp_path / "lucy6.ptbs")
    sc.write_ptbs(path, sc.lucy_standin(sc.read_ptbs(cornel_ptbs), level=6))
    W, H = 160, 90
    o, pt = _pair(ptb, path, W, H)
    o.configure(W, H, MODE_BVH, nee=0, max_depth=6)
    _, _, rrng, st = o.render(1)
    pt.set_params(width=W, height=H, mode=MODE_BVH, nee=0, max_depth=6, count_tests=1)
    pt.reset()
    pt.render(1)
    c = pt.counters()
    _, rng = pt.read_state()
    # explicit rays through ptb_trace_rays: same tree, this call's counters
    rays = sc.soup_rays(1 << 12)
    t1, tri1 = pt.trace_rays(rays, MODE_BVH)
    pt.use_ref_bvh(False)
    t0, tri0 = pt.trace_rays(rays, MODE_BVH)
    pt.close()
    o.close()
    assert (c["tri_tests"], c["as_tests"]) == (st["tri_tests"], st["as_tests"])
    assert int((rng != rrng).sum()) <= 2
    assert (_bits(t0) == _bits(t1)).all() and (tri0 == tri1).all()


def test_upload_validation(ptb, cornel_ptbs):
    o, pt = _pair(ptb, cornel_ptbs, 8, 8)
    nodes, idx, roots, root = o.export_bvh()
    o.close()
    n = nodes.view(ptb.REF_BVH_NODE).copy()
    bad = n.copy()
    bad["child_left"][root] = len(n) + 5
    with pytest.raises(ptb.PtbError):
        pt.upload_ref_bvh(bad, idx, roots, root)
    cyc = n.copy()
    inner = int(np.nonzero(cyc["child_left"] >= 0)[0][-1])
    cyc["child_left"][inner] = root  # a link back to the root
    with pytest.raises(ptb.PtbError):
        pt.upload_ref_bvh(cyc, idx, roots, root)
    with pytest.raises(ptb.PtbError):
        pt.upload_ref_bvh(n, idx, roots, len(n))
    with pytest.raises(ptb.PtbError):  # a failed upload leaves no tree behind
        pt.use_ref_bvh(True)
    pt.upload_ref_bvh(n, idx, roots, root)
    pt.use_ref_bvh(True)
    pt.close()


@pytest.mark.skipif(not (have_facade() and have_ref()), reason="needs oracle/_ref/ptfacade and ptref (built where the reference sources are)")
def test_facade_reports_the_cpu_tracers_counts(cornel_ptbs, ptb):
    """Drop-in flavour: the reference's own objects + host/tracer_facade.cpp.  The facade reads the tree the reference's
    constructBVH built through getNode / getNodeAABB / bvhGetTrisIndex and must then report exactly the counts the
    UNMODIFIED reference reports for the same render."""
    W, H, spp = 96, 64, 2
    _, ref, _ = ptref_render(cornel_ptbs, W, H, MODE_BVH, spp, max_depth=0, nee=0, want_image=False)
    img, rec = ptfacade_render(cornel_ptbs, W, H, MODE_BVH, spp, 0, 0, extra=("--ref-tree", 1))
    assert (rec["tri_tests"], rec["as_tests"]) == (ref["tri_tests"], ref["as_tests"])
    _, own = ptfacade_render(cornel_ptbs, W, H, MODE_BVH, spp, 0, 0)
    assert own["as_tests"] != ref["as_tests"]  # without the tree the counts are those of the 64-byte SAH nodes
