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

import numpy as np
import pytest

from oracle_lib import MODE_ALL, MODE_BVH, MODE_SUB_BEAMS, MODE_TWO_PLANE, Oracle, have_ref, ptref_primary_hits, ptref_render

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
