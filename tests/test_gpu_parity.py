"""GPU parity: the sm_100a path (through the C ABI, libptb.so) against the CPU oracle
(oracle/libptoracle.so, pinned to the compiled reference by tests/test_oracle_pins.py) on the
same scene, camera and RNG seeds.

Bars (BASELINE.json north_star):
  * primary-ray hit triangle ids equal, except documented ties (bit-equal t on two triangles),
    where t must agree within 1e-5 relative — here t is required to be BIT-equal everywhere;
  * images: the arithmetic is replayed operation for operation (no FMA, same libm for the
    per-material pow), so images are required to match to <= 2e-6 relative per pixel, with the
    number of pixels that differ at all reported.
"""
import numpy as np
import pytest

from oracle_lib import MODE_ALL, MODE_BVH, MODE_SUB_BEAMS, MODE_TWO_PLANE, Oracle

pytestmark = pytest.mark.gpu

W = H = 128
GS, OS, M = 10, 5, 40


@pytest.fixture(scope="module")
def oracle(cornel_ptbs):
    o = Oracle(cornel_ptbs)
    o.build(grid=True, beams=True, gs=GS, os_=OS, m=M, bgs=GS, bm=M)
    yield o
    o.close()


@pytest.fixture(scope="module")
def gpu(ptb, cornel_ptbs):
    pt = ptb.PathTracer(0)
    pt.load_ptbs(cornel_ptbs)
    pt.build_bvh()
    pt.build_subgrid(GS, OS, M)
    pt.build_subbeams(GS, M)
    yield pt
    pt.close()


def _hits_agree(t, tri, rt, rtri):
    """ids equal except ties; returns (#tie pixels, #bad pixels)."""
    same_t = t.view(np.uint32) == rt.view(np.uint32)
    bad = ~same_t
    ties = same_t & (tri != rtri)
    return int(ties.sum()), int(bad.sum())


@pytest.mark.parametrize("mode", [MODE_ALL, MODE_BVH, MODE_TWO_PLANE, MODE_SUB_BEAMS])
def test_primary_hits_cornell(gpu, oracle, mode):
    gpu.set_params(width=W, height=H, mode=mode, nee=0, max_depth=0)
    t, tri, mat = gpu.primary_hits()
    oracle.configure(W, H, mode)
    rt, rtri, rmat = oracle.primary_hits()
    ties, bad = _hits_agree(t, tri, rt, rtri)
    assert bad == 0, f"{bad} primary rays with a different hit distance"
    agree = float((tri == rtri).mean())
    assert agree >= 0.9999 or ties == int((tri != rtri).sum())
    # ties keep the same material on Cornell (coplanar quads of one object)
    assert int(((mat != rmat) & (tri == rtri)).sum()) == 0


def _render_both(gpu, oracle, mode, spp, nee, max_depth):
    gpu.set_params(width=W, height=H, mode=mode, nee=nee, max_depth=max_depth, count_tests=0)
    gpu.reset()
    gpu.render(spp)
    img = gpu.read_image()
    samples, rng = gpu.read_state()
    oracle.configure(W, H, mode, nee=nee, max_depth=max_depth)
    rimg, rs, rrng, st = oracle.render(spp)
    return img, samples, rng, rimg, rs, rrng, st


def _img_close(img, rimg):
    diff = np.abs(img[..., :3] - rimg[..., :3])
    scale = np.maximum(np.abs(rimg[..., :3]), 1e-3)
    rel = float((diff / scale).max())
    nbits = int((img.view(np.uint32) != rimg.view(np.uint32)).any(axis=-1).sum())
    return rel, nbits


@pytest.mark.parametrize("mode", [MODE_BVH, MODE_TWO_PLANE, MODE_SUB_BEAMS, MODE_ALL])
@pytest.mark.parametrize("nee", [0, 1])
def test_render_unbounded_matches_oracle(gpu, oracle, mode, nee):
    img, samples, rng, rimg, rs, rrng, st = _render_both(gpu, oracle, mode, spp=3, nee=nee, max_depth=0)
    assert (samples == rs).all()
    stream_diff = int((rng != rrng).sum())
    rel, nbits = _img_close(img, rimg)
    print(f"mode {mode} nee {nee}: rng streams differing {stream_diff}, pixels differing {nbits}, max rel {rel:.3g}")
    # the RNG stream after 3 samples is the strictest check: every branch of every bounce agreed
    assert stream_diff <= max(1, W * H // 2000)
    assert rel <= 2e-6 or nbits <= max(1, W * H // 2000)
    c = gpu.counters()
    assert c["paths"] == st["paths"]


@pytest.mark.parametrize("max_depth", [2, 5])
def test_render_depth_capped(gpu, oracle, max_depth):
    """BASELINE configs 1 and 2: primary + 1 bounce, primary + 4 bounces."""
    img, samples, rng, rimg, rs, rrng, st = _render_both(gpu, oracle, MODE_BVH, spp=4, nee=1, max_depth=max_depth)
    assert (samples == rs).all()
    assert int((rng != rrng).sum()) <= max(1, W * H // 2000)
    rel, nbits = _img_close(img, rimg)
    assert rel <= 2e-6 or nbits <= max(1, W * H // 2000)


def test_progressive_equals_one_shot(gpu):
    """Raising the sample target continues the per-pixel streams (tracer.cpp:148; window.cpp:289)."""
    gpu.set_params(width=W, height=H, mode=MODE_BVH, nee=0, max_depth=0)
    gpu.reset()
    gpu.render(4)
    a = gpu.read_image()
    gpu.reset()
    gpu.render(1)
    gpu.render(3)
    b = gpu.read_image()
    assert (a.view(np.uint32) == b.view(np.uint32)).all()


def test_trace_rays_matches_primary(gpu, oracle):
    oracle.configure(W, H, MODE_BVH)
    rt, rtri, rmat, rays = oracle.primary_hits(want_rays=True)
    for mode in (MODE_ALL, MODE_BVH, MODE_TWO_PLANE, MODE_SUB_BEAMS):
        t, tri = gpu.trace_rays(rays, mode)
        assert (t.view(np.uint32) == rt.view(np.uint32)).all(), f"mode {mode}"


def test_tile_partition_is_bitwise_identical(ptb, cornel_ptbs, gpu):
    """Multi-GPU correctness on one device: the union of the ranks' pixel sets, rendered by
    separate contexts, equals the single-context frame bit for bit."""
    gpu.set_params(width=W, height=H, mode=MODE_BVH, nee=1, max_depth=0, rank=0, world=1)
    gpu.reset()
    gpu.render(2)
    full = gpu.read_image()
    acc = np.zeros_like(full)
    world = 3
    for r in range(world):
        pt = ptb.PathTracer(0)
        pt.load_ptbs(cornel_ptbs)
        pt.build_bvh()
        pt.set_params(width=W, height=H, mode=MODE_BVH, nee=1, max_depth=0, rank=r, world=world, tile_w=16, tile_h=4)
        pt.reset()
        pt.render(2)
        acc += pt.read_image()  # non-owned pixels are exactly zero -> the sum is exact
        pt.close()
    assert (acc.view(np.uint32) == full.view(np.uint32)).all()


def test_heat_map_view(gpu):
    """Debug view (tracer.cpp:136-142): one ray per pixel, pixel = test count / scale.  In mode ALL every
    ray tests every triangle (object.cpp:19-34), so the triangle heat map is flat at n_tris / scale; in
    BVH mode the heat map must sum to the counters the same pass reports."""
    scale = 404
    gpu.set_params(width=W, height=H, mode=MODE_ALL, nee=0, max_depth=0, debug_view=2, debug_scale=scale, count_tests=1)
    gpu.reset()
    gpu.render(1)
    img = gpu.read_image()
    assert (img[..., 0] == np.float32(202) / np.float32(scale)).all() and (img[..., 1] == img[..., 0]).all()
    scale = 1 << 20
    gpu.set_params(mode=MODE_BVH, debug_view=1, debug_scale=scale)
    gpu.reset()
    gpu.render(1)
    img = gpu.read_image()
    s, _ = gpu.read_state()
    c = gpu.counters()
    assert (s == 1).all() and c["paths"] == W * H
    assert int(np.rint(img[..., 0].astype(np.float64) * scale).sum()) == c["as_tests"]
    over = 3
    gpu.set_params(debug_view=1, debug_scale=over)
    gpu.reset()
    gpu.render(1)
    red = gpu.read_image()
    assert ((red[..., 0] == 1.0) & (red[..., 1] == 0.0)).any()  # count > scale -> red
    gpu.set_params(debug_view=0, debug_scale=0, count_tests=0)


def test_clear_keeps_rng_streams(gpu, oracle):
    """reset() (tracer.cpp:195-217) zeroes image and sample counts but does not reseed: the next render
    continues every pixel's stream, exactly like the oracle resumed from zeroed image + kept rng."""
    gpu.set_params(width=W, height=H, mode=MODE_BVH, nee=0, max_depth=4)
    gpu.reset()
    gpu.render(2)
    _, rng_a = gpu.read_state()
    gpu.clear()
    s, rng_b = gpu.read_state()
    assert (s == 0).all() and (rng_a == rng_b).all() and (gpu.read_image() == 0).all()
    gpu.render(1)
    img = gpu.read_image()
    oracle.configure(W, H, MODE_BVH, nee=0, max_depth=4)
    _, _, r2, _ = oracle.render(2)
    rimg, _, _, _ = oracle.render(1, state=(np.zeros((H, W, 4), np.float32), np.zeros(W * H, np.int32), r2.copy()))
    rel, nbits = _img_close(img, rimg)
    assert rel <= 2e-6 or nbits <= max(1, W * H // 2000)


def test_camera_basis_equals_scene_camera(gpu, ptb, cornel_ptbs):
    gpu.set_params(width=W, height=H, mode=MODE_BVH, nee=0, max_depth=3)
    gpu.reset()
    gpu.render(1)
    a = gpu.read_image()
    cam = ptb.scenes.read_ptbs(cornel_ptbs)["camera"]
    # cornel.scene looks down +z: orthoNormalized((0,0,1),(0,1,0),(-1,0,0)) is the identity basis
    import ctypes
    libm = ctypes.CDLL("libm.so.6")
    libm.tanf.restype = ctypes.c_float
    libm.tanf.argtypes = [ctypes.c_float]
    fov_scale = libm.tanf(ctypes.c_float(float(np.float32(np.float32(70 * np.float32(3.14)) / np.float32(180.0)) / np.float32(2.0))))
    gpu.set_camera_basis(cam["pos"], [0, 0, 1], [-1, 0, 0], [0, 1, 0], cam["focus"], cam["dof"], fov_scale)
    gpu.reset()
    gpu.render(1)
    b = gpu.read_image()
    gpu.set_camera(cam["pos"], cam["forward"], cam["focus"], cam["dof"])
    assert (a.view(np.uint32) == b.view(np.uint32)).all()


def test_counters_match_oracle_ray_counts(gpu, oracle):
    for mode in (MODE_BVH, MODE_TWO_PLANE):
        gpu.set_params(width=W, height=H, mode=mode, nee=0, max_depth=0)
        gpu.reset()
        gpu.render(2)
        c = gpu.counters()
        oracle.configure(W, H, mode, nee=0, max_depth=0)
        _, _, _, st = oracle.render(2)
        assert c["rays"] == st["rays_unique"], f"mode {mode}: GPU traversals vs oracle unique rays"
        assert c["ref_rays"] == st["rays_traversed"], f"mode {mode}: replayed reference ray count"


def test_sample_sharded_frame_is_statistically_equivalent(gpu, ptb):
    """Opt-in sample-range sharding (multigpu.combine_sample_shards): 4 shards x 16 spp from per-shard seeds
    against single 64-spp frames from three other seed sets and a 2048-spp reference.  Stated bar: per-pixel
    RMSE of linear radiance (clamped to 4 to keep single fireflies from deciding the test) against the
    reference no worse than 1.15x the worst of the single runs — same estimator, different random numbers."""
    mg = ptb.multigpu
    Wd = Hd = 96
    spp, shards = 64, 4

    def frame(n, seed_offset):
        gpu.set_params(width=Wd, height=Hd, mode=MODE_BVH, nee=1, max_depth=5, seed_offset=seed_offset)
        gpu.reset()
        gpu.render(n)
        return gpu.read_image().astype(np.float64)[..., :3] * (n + 1) / n  # undo the N+1 law

    def rmse(a, b):
        return float(np.sqrt(((np.minimum(a, 4.0) - np.minimum(b, 4.0)) ** 2).mean()))

    ref = frame(2048, 99)
    singles = [rmse(frame(spp, s), ref) for s in (0, 21, 22)]
    total = np.zeros_like(ref)
    for r in range(shards):
        n = mg.sample_share(spp, r, shards)
        total += frame(n, r) * n
    sharded = total / spp
    gpu.set_params(seed_offset=0)
    r_sh = rmse(sharded, ref)
    print(f"RMSE vs 2048 spp: single 64-spp runs {singles}, 4 sample shards {r_sh:.4f}")
    assert r_sh <= 1.15 * max(singles)
    assert abs(sharded.mean() - ref.mean()) <= 0.03 * ref.mean()


def test_schedulers_give_identical_frames(ptb, cornel_ptbs, monkeypatch):
    """The path kernel (PTB_MEGA=1, persistent lanes) and the wavefront pipeline (0) share the arithmetic and differ only in
    scheduling: frames, RNG states and counters must be bit-identical.  (The two rejected schedulers of round 1, k_duo = 3 and
    k_pool = 2, are only in builds with -DPTB_WITH_ALT_SCHEDULERS; PTB_TEST_ALT_SCHEDULERS=1 adds them here.)"""
    import os
    outs = []
    for sched in ("1", "3", "2", "0") if os.environ.get("PTB_TEST_ALT_SCHEDULERS") else ("1", "0"):
        monkeypatch.setenv("PTB_MEGA", sched)
        pt = ptb.PathTracer(0)
        pt.load_ptbs(cornel_ptbs)
        pt.build_bvh()
        pt.build_subgrid(GS, OS, M)
        res = []
        for mode, nee, md in ((MODE_BVH, 1, 0), (MODE_BVH, 0, 5), (MODE_TWO_PLANE, 1, 4)):
            pt.set_params(width=W, height=H, mode=mode, nee=nee, max_depth=md)
            pt.reset()
            pt.render(2)
            pt.render(1)
            s, r = pt.read_state()
            c = pt.counters()
            res.append((pt.read_image().copy(), s, r, c["rays"], c["paths"], c["ref_rays"], c["shadow_rays"]))
        pt.close()
        outs.append(res)
    for other in outs[1:]:
        for a, b in zip(outs[0], other):
            assert (a[0].view(np.uint32) == b[0].view(np.uint32)).all()
            assert (a[1] == b[1]).all() and (a[2] == b[2]).all()
            assert a[3:] == b[3:]


def test_async_render_and_edge_cases(ptb, cornel_ptbs, gpu):
    """ptb_render_async + ptb_sync equals ptb_render; reads settle a pending render; ragged frame sizes;
    call-order errors are reported, not crashed on."""
    pt = ptb.PathTracer(0)
    with pytest.raises(ptb.PtbError):  # nothing uploaded / no params
        pt.render(1)
    pt.load_ptbs(cornel_ptbs)
    pt.set_params(width=50, height=37, mode=MODE_BVH, nee=1, max_depth=3)  # not multiples of 4 or 32
    with pytest.raises(ptb.PtbError) as e:
        pt.render(1)
    assert "ptb_build_bvh" in str(e.value)
    pt.build_bvh()
    pt.set_params(mode=MODE_TWO_PLANE)
    with pytest.raises(ptb.PtbError):
        pt.render(1)  # Subgrid not built
    with pytest.raises(ptb.PtbError):
        pt.set_params(width=0)
    pt.set_params(width=50, height=37, mode=MODE_BVH)
    pt.reset()
    pt.render(3)
    a = pt.read_image()
    pt.reset()
    pt.render_async(2)
    pt.render_async(1)   # settles the first, enqueues the second
    b = pt.read_image()  # settles the second
    pt.sync()
    assert (a.view(np.uint32) == b.view(np.uint32)).all()
    assert pt.counters()["paths"] == 50 * 37 * 3
    gpu.set_params(width=50, height=37, mode=MODE_BVH, nee=1, max_depth=3)
    gpu.reset()
    gpu.render(3)
    assert (gpu.read_image().view(np.uint32) == a.view(np.uint32)).all()
    # an empty scene renders black and never touches a structure
    pt2 = ptb.PathTracer(0)
    pt2.upload_triangles(np.zeros((0, 9), np.float32), np.zeros((0, 9), np.float32), None, np.zeros(0, np.int32))
    pt2.upload_materials(b"", 0)
    pt2.upload_objects(b"", 0)
    pt2.set_camera([0, 0, -5], [0, 0, 1])
    pt2.build_bvh()
    pt2.set_params(width=16, height=8, mode=MODE_BVH, max_depth=4)
    pt2.reset()
    pt2.render(2)
    img = pt2.read_image()
    assert (img[..., :3] == 0).all() and (pt2.read_state()[0] == 2).all()
    pt2.close()
    pt.close()


def test_full_size_config2_properties(ptb, cornel_ptbs, gpu):
    """BASELINE config 2 at its full size (1024x1024, 64 spp, max_depth 5, NEE off) through properties that do not
    need the CPU oracle to finish: exact path / sample / ray bookkeeping, finite non-negative radiance, the
    one-shot frame equal to 16+48 progressive samples and to the union of 8 tile-sharded ranks.  (Bit equality
    with the oracle for the same configuration at 128x128 is test_render_depth_capped[5].)"""
    Wf = Hf = 1024
    spp = 64
    gpu.set_params(width=Wf, height=Hf, mode=MODE_BVH, nee=0, max_depth=5, rank=0, world=1)
    gpu.reset()
    gpu.render(spp)
    full = gpu.read_image()
    samples, rng = gpu.read_state()
    c = gpu.counters()
    assert (samples == spp).all()
    assert c["paths"] == Wf * Hf * spp
    assert Wf * Hf * spp < c["rays"] <= 5 * Wf * Hf * spp  # primary + at most 4 bounces
    assert c["ref_rays"] >= c["rays"]                        # the reference re-traces misses
    assert np.isfinite(full).all() and (full[..., :3] >= 0).all() and (full[..., 3] == 1).all()
    assert 0.05 < full[..., :3].mean() < 5.0
    # progressive == one shot (the RNG stream and the N+1 law carry across ptb_render calls)
    gpu.reset()
    gpu.render(16)
    gpu.render(48)
    again = gpu.read_image()
    assert (again.view(np.uint32) == full.view(np.uint32)).all()
    assert (gpu.read_state()[1] == rng).all()
    # 8 ranks' tiles (the 8-GPU partition, rendered one after another on this device) == the frame
    acc = np.zeros_like(full)
    for r in range(8):
        gpu.set_params(rank=r, world=8, tile_w=32, tile_h=8)
        gpu.reset()
        gpu.render(spp)
        acc += gpu.read_image()
    gpu.set_params(rank=0, world=1)
    assert (acc.view(np.uint32) == full.view(np.uint32)).all()
