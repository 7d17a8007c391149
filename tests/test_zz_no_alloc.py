"""ptb_trace_rays_static — the entry point that allocates nothing (scratch in a file-scope `static __device__` array of
libptb.so; include/ptb.h) — against the existing ptb_trace_rays, through the C ABI, on two calls in a row.

Collected last on purpose.  Passes alone in a fresh process and after the other tests: it
creates its own context and depends on no state they leave.  Tolerance: the same as every traversal test here — hit
distances bit-equal, triangle ids equal (same kernel, same tree, so not even ties may differ)."""
import numpy as np
import pytest

from oracle_lib import MODE_BVH

pytestmark = pytest.mark.gpu


def _rays(n, seed):
    """incoherent rays towards the Cornell room (it spans roughly [-280, 280]^3 around the origin)"""
    rng = np.random.default_rng(seed)
    o = rng.normal(size=(n, 3))
    o = 900.0 * o / np.linalg.norm(o, axis=1, keepdims=True)
    target = rng.uniform(-250.0, 250.0, size=(n, 3))
    d = target - o
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return np.concatenate([o, d], axis=1).astype(np.float32)


def test_static_scratch_entry_point_equals_trace_rays_twice_in_a_row(ptb, cornel_ptbs):
    pt = ptb.PathTracer(0)
    pt.load_ptbs(cornel_ptbs)
    pt.build_bvh()
    pt.set_params(width=64, height=64, mode=MODE_BVH, nee=0, max_depth=0)
    limit = int(pt.L.ptb_trace_rays_static_max())
    assert limit == 65536
    # the second call is SHORTER and over other rays: whatever the first one left in the shared array must not show
    for n, seed in ((limit, 11), (limit // 3 + 5, 12)):
        od = _rays(n, seed)
        t_ref, tri_ref = pt.trace_rays(od, MODE_BVH)
        t, tri = pt.trace_rays_static(od)
        assert (t_ref < np.inf).mean() > 0.5, "the test rays should mostly hit the room"
        assert (t.view(np.uint32) == t_ref.view(np.uint32)).all()
        assert (tri == tri_ref).all()
    # a second context on the same device shares the array (sequential use is safe)
    pt2 = ptb.PathTracer(0)
    pt2.load_ptbs(cornel_ptbs)
    pt2.build_bvh()
    od = _rays(1000, 13)
    t_ref, tri_ref = pt.trace_rays(od, MODE_BVH)
    t, tri = pt2.trace_rays_static(od)
    assert (t.view(np.uint32) == t_ref.view(np.uint32)).all() and (tri == tri_ref).all()
    pt2.close()
    # beyond the maximum: a clear error, nothing launched
    with pytest.raises(ptb.PtbError, match="exceed the static scratch"):
        pt.trace_rays_static(_rays(limit + 1, 14))
    t, tri = pt.trace_rays_static(np.zeros((0, 6), np.float32))  # empty batch
    assert t.shape == (0,) and tri.shape == (0,)
    pt.close()


def test_static_entry_point_needs_a_built_tree(ptb, cornel_ptbs):
    pt = ptb.PathTracer(0)
    pt.load_ptbs(cornel_ptbs)
    with pytest.raises(ptb.PtbError, match="ptb_build_bvh"):
        pt.trace_rays_static(_rays(16, 1))
    pt.close()
