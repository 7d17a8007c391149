"""ptb_group_*: the frame tile-sharded over several contexts behind one host thread, gathered on member 0 by one kernel
that reads every pixel from its owner (csrc/ptb_group.cu).  The bar is the multi-GPU correctness property of SURVEY.md
8e: the gathered frame equals the 1-GPU frame bit for bit.  A device may be listed twice, so the tile rule, the
concurrent enqueue and the gather kernel are exercised on a single GPU too; with >= 2 GPUs the peers are real."""
import numpy as np
import pytest

from oracle_lib import MODE_BVH, MODE_TWO_PLANE

pytestmark = pytest.mark.gpu


def _bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def _single(ptb, path, W, H, spp, **kw):
    pt = ptb.PathTracer(0)
    pt.load_ptbs(path)
    pt.build_bvh()
    pt.build_subgrid(10, 5, 40, gpu=True)
    pt.set_params(width=W, height=H, **kw)
    pt.reset()
    pt.render(spp)
    out = pt.read_image(), pt.counters()
    pt.close()
    return out


def _group(ptb, devices, path):
    g = ptb.PathTracerGroup(devices)

    def prep(m):
        m.load_ptbs(path)
        m.build_bvh()
        m.build_subgrid(10, 5, 40, gpu=True)
    g.each(prep)
    return g


@pytest.mark.parametrize("devices,W,H", [([0, 0], 160, 120), ([0, 0, 0], 100, 52), ([0], 64, 48)])
def test_group_frame_is_the_single_gpu_frame(ptb, cornel_ptbs, devices, W, H):
    spp = 3
    g = _group(ptb, devices, cornel_ptbs)
    assert all(g.peer_direct())
    for mode, nee, depth in ((MODE_BVH, 1, 0), (MODE_TWO_PLANE, 0, 4)):
        ref, rc = _single(ptb, cornel_ptbs, W, H, spp, mode=mode, nee=nee, max_depth=depth)
        g.set_params(width=W, height=H, mode=mode, nee=nee, max_depth=depth)
        g.reset()
        ms = g.render(2)
        g.render(spp - 2)  # progressive: the members continue their own pixels' streams
        img = g.read_image()
        c = g.counters()
        assert ms > 0.0
        assert (_bits(img) == _bits(ref)).all(), f"{len(devices)} members, mode {mode}: gathered frame differs from the 1-GPU frame"
        assert c["paths"] == rc["paths"] == W * H * spp and c["rays"] == rc["rays"]
    # every member rendered only its tiles
    mask = ptb.multigpu.owned_mask(W, H, 0, len(devices))
    own = g.members[0].read_image()
    assert (own[~mask][:, :3] == 0).all() and (len(devices) == 1 or (own[mask][:, :3] != 0).any())
    g.clear()
    assert (g.read_image()[..., :3] == 0).all()
    g.close()


def test_group_errors_are_loud(ptb, cornel_ptbs):
    g = ptb.PathTracerGroup([0, 0])
    with pytest.raises(ptb.PtbError):
        g.render(1)  # no params
    g.set_params(width=32, height=32, mode=MODE_BVH)
    with pytest.raises(ptb.PtbError):
        g.render(1)  # no scene / structure on the members
    g.close()
    with pytest.raises(ptb.PtbError):
        ptb.PathTracerGroup([0, 4096])


def test_group_over_real_peers(ptb, cornel_ptbs):
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("one GPU on this box")
    sc = ptb.scenes
    scene = sc.lucy_standin(sc.read_ptbs(cornel_ptbs), level=7)
    W, H, spp = 1920, 1080, 8
    pt = ptb.PathTracer(0)
    sc.upload_scene(pt, scene)
    pt.build_bvh()
    pt.set_params(width=W, height=H, mode=MODE_BVH, nee=0, max_depth=9)
    pt.reset()
    t1 = pt.render(spp)
    ref = pt.read_image()
    pt.close()
    g = ptb.PathTracerGroup(list(range(n)))

    def prep(m):
        sc.upload_scene(m, scene)
        m.build_bvh()
    g.each(prep)
    g.set_params(width=W, height=H, mode=MODE_BVH, nee=0, max_depth=9)
    g.reset()
    tn = g.render(spp)
    img = g.read_image()
    print(f"{n} GPUs in one process: {t1:.2f} ms -> {tn:.2f} ms per {spp}-spp 1080p frame ({t1 / tn:.2f}x), peers read in place: {g.peer_direct()}")
    g.close()
    assert (_bits(img) == _bits(ref)).all()


def test_group_sample_sharding_is_the_documented_combination(ptb, cornel_ptbs):
    """PTB_SHARD_SAMPLES: member i renders all pixels with its share of the samples from seeds offset by i; the combined
    frame is sum_i(mean_i * (n_i + 1)) / (n + 1) — checked against the same arithmetic on frames of single contexts."""
    W, H, spp = 96, 64, 7
    g = _group(ptb, [0, 0, 0], cornel_ptbs)
    g.set_sharding(True)
    g.set_params(width=W, height=H, mode=MODE_BVH, nee=1, max_depth=5)
    g.reset()
    g.render(4)
    g.render(spp - 4)  # shares per call: 4 -> (2,1,1), 3 -> (1,1,1)
    img = g.read_image()
    assert g.counters()["paths"] == W * H * spp
    g.close()
    acc = np.zeros((H, W, 3), np.float32)
    for i, n_i in enumerate((3, 2, 2)):
        pt = ptb.PathTracer(0)
        pt.load_ptbs(cornel_ptbs)
        pt.build_bvh()
        pt.set_params(width=W, height=H, mode=MODE_BVH, nee=1, max_depth=5, seed_offset=i)
        pt.reset()
        pt.render(n_i)
        acc = acc + pt.read_image()[..., :3] * np.float32(n_i + 1)
        pt.close()
    want = acc / np.float32(spp + 1)
    assert (_bits(img[..., :3]) == _bits(want.astype(np.float32))).all()
    assert (img[..., 3] == 1.0).all()
