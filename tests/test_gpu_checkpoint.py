"""Row n4: the reference's progressive state is (image, samples[x][y], randomStates[x][y]) (src/tracer.cpp:24-26,148).
ptb_read_image + ptb_read_state dump it, ptb_write_state restores it into a NEW context, and the continued render must equal
the uninterrupted one bit for bit — also from a NON-uniform state (pixels at different sample counts), which switches the path
kernel to its one-item-per-pixel mode (check_ready = 0, csrc/ptb_api.cu render_mega)."""
import numpy as np
import pytest

from oracle_lib import MODE_BVH, MODE_TWO_PLANE

pytestmark = pytest.mark.gpu


def _ctx(ptb, cornel_ptbs, W, H, mode, nee, depth):
    pt = ptb.PathTracer(0)
    pt.load_ptbs(cornel_ptbs)
    pt.build_bvh()
    if mode == MODE_TWO_PLANE:
        pt.build_subgrid(10, 5, 40)
    pt.set_params(width=W, height=H, mode=mode, nee=nee, max_depth=depth)
    pt.reset()
    return pt


def _same(a, b):
    return bool((a[0].view(np.uint32) == b[0].view(np.uint32)).all() and (a[1] == b[1]).all() and (a[2] == b[2]).all())


def _dump(pt):
    s, r = pt.read_state()
    return pt.read_image().copy(), s.copy(), r.copy()


@pytest.mark.parametrize("W,H,mode,nee,depth", [(160, 96, MODE_BVH, 1, 0), (1024, 768, MODE_BVH, 0, 5), (96, 64, MODE_TWO_PLANE, 1, 4)])
def test_dump_destroy_restore_continue_equals_uninterrupted(ptb, cornel_ptbs, W, H, mode, nee, depth):
    k, m = 3, 4
    pt = _ctx(ptb, cornel_ptbs, W, H, mode, nee, depth)
    pt.render(k + m)
    whole = _dump(pt)
    pt.close()

    pt = _ctx(ptb, cornel_ptbs, W, H, mode, nee, depth)
    pt.render(k)
    part = _dump(pt)
    pt.close()  # destroy: nothing of the first context survives

    pt = _ctx(ptb, cornel_ptbs, W, H, mode, nee, depth)
    pt.write_state(*part)
    assert _same(_dump(pt), part), "write -> read must round-trip"
    pt.render(m)
    resumed = _dump(pt)
    assert (resumed[1] == k + m).all()
    assert _same(resumed, whole), "restored + continued render differs from the uninterrupted one"

    # non-uniform state: even columns as after k samples, odd columns as after k + 2
    pt.reset()
    pt.render(k + 2)
    later = _dump(pt)
    mixed = [a.copy() for a in part]
    cols = np.arange(W) % 2 == 1
    mixed[0][:, cols] = later[0][:, cols]
    for i in (1, 2):
        a, b = mixed[i].reshape(H, W), later[i].reshape(H, W)
        a[:, cols] = b[:, cols]
        mixed[i] = a.reshape(-1)
    pt.write_state(*mixed)
    pt.render(m - 2)  # target = max(samples) + (m - 2) = k + m for every pixel
    resumed = _dump(pt)
    pt.close()
    assert (resumed[1] == k + m).all()
    assert _same(resumed, whole), "continuing a non-uniform state differs from the uninterrupted render"
