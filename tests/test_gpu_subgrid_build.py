"""The Subgrid ("2Plane"/Hybrid) structure with its channels filled on the GPU (ptb_build_subgrid_gpu,
csrc/ptb_grid_gpu.cu) against the host builder: both evaluate the same separating-axis code
(csrc/ptb_sat.h), so the bar is byte equality of the three device pools (ptb_structure_digest), plus the usual
closest-hit parity of the structure that comes out."""
import numpy as np
import pytest

from oracle_lib import MODE_ALL, MODE_BVH, MODE_TWO_PLANE, Oracle

pytestmark = pytest.mark.gpu


def _bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def _digests(pt, gs, os_, m):
    pt.build_subgrid(gs, os_, m, gpu=False)
    host = (pt.structure_digest(MODE_TWO_PLANE), pt.structure_bytes(MODE_TWO_PLANE))
    wall, dev = pt.build_subgrid(gs, os_, m, gpu=True)
    gpu = (pt.structure_digest(MODE_TWO_PLANE), pt.structure_bytes(MODE_TWO_PLANE))
    assert host[0] != 0 and gpu[0] != 0
    return host, gpu, wall, dev


@pytest.mark.parametrize("gs,os_,m", [(10, 5, 40), (8, 8, 50), (4, 3, 100), (3, 2, 500)])
def test_cornell_pools_identical_to_host_builder(ptb, cornel_ptbs, gs, os_, m):
    """m = 100 / 500: leaf grids with more than 64 candidates (pass 2 repeats the tests instead of expanding masks);
    m = 500 > 202 triangles is the flat classic two-plane grid."""
    pt = ptb.PathTracer(0)
    pt.load_ptbs(cornel_ptbs)
    pt.build_bvh()
    host, gpu, wall, dev = _digests(pt, gs, os_, m)
    assert host == gpu, f"pools differ: host {host}, gpu {gpu}"
    # and the structure renders the oracle's frame
    W = H = 128
    o = Oracle(cornel_ptbs)
    o.build(grid=True, gs=gs, os_=os_, m=m)
    o.configure(W, H, MODE_TWO_PLANE, nee=1, max_depth=0)
    rimg, rs, rrng, _ = o.render(2)
    o.close()
    pt.set_params(width=W, height=H, mode=MODE_TWO_PLANE, nee=1, max_depth=0)
    pt.reset()
    pt.render(2)
    s, rng = pt.read_state()
    img = pt.read_image()
    pt.close()
    assert (s == rs).all() and int((rng != rrng).sum()) <= W * H // 2000
    assert int((_bits(img) != _bits(rimg)).any(axis=-1).sum()) <= W * H // 2000


def test_lucy_level6_and_soup_pools_identical(ptb, cornel_ptbs):
    sc = ptb.scenes
    for name, scene, (gs, os_, m) in (("lucy level 6", sc.lucy_standin(sc.read_ptbs(cornel_ptbs), level=6), (6, 4, 50)),
                                     ("soup 1e4", sc.triangle_soup(10_000), (6, 4, 40))):
        pt = ptb.PathTracer(0)
        sc.upload_scene(pt, scene)
        pt.build_bvh()
        host, gpu, wall, dev = _digests(pt, gs, os_, m)
        print(f"{name}: {gpu[1] / 1e6:.1f} MB, GPU build {wall:.1f} ms wall / {dev:.1f} ms device")
        assert host == gpu, f"{name}: pools differ"
        pt.close()


def test_empty_scene_and_device_tree(ptb, cornel_ptbs):
    sc = ptb.scenes
    pt = ptb.PathTracer(0)
    sc.upload_scene(pt, sc.scene_from_triangles(np.zeros((0, 9), np.float32)))
    pt.build_bvh()
    host, gpu, _, _ = _digests(pt, 4, 4, 40)
    assert host == gpu
    pt.set_params(width=16, height=8, mode=MODE_TWO_PLANE)
    t, tri, _ = pt.primary_hits()
    assert np.isinf(t).all()
    pt.close()
    # both builders on the device: BVH from ptb_build_bvh_gpu, channels from ptb_build_subgrid_gpu
    pt = ptb.PathTracer(0)
    pt.load_ptbs(cornel_ptbs)
    pt.build_bvh(gpu=True)
    host, gpu, _, _ = _digests(pt, 10, 5, 40)  # the host builder walks a mirror of the device tree
    assert host == gpu
    W = H = 160
    pt.set_params(width=W, height=H, mode=MODE_TWO_PLANE)
    t2, tri2, _ = pt.primary_hits()
    pt.set_params(mode=MODE_ALL)
    ta, tria, _ = pt.primary_hits()
    pt.close()
    assert int((_bits(t2) != _bits(ta)).sum()) == 0 and float((tri2 == tria).mean()) >= 0.9999


def test_lucy_full_size_gpu_built_subgrid(ptb, cornel_ptbs):
    """BASELINE config 3's Subgrid (gs 8 / os 8 / m 50, the published parameters) for the 524,310-triangle stand-in,
    built on the GPU: the same closest hits as the BVH at 1920x1080 (SURVEY.md App. B.2 property)."""
    sc = ptb.scenes
    scene = sc.lucy_standin(sc.read_ptbs(cornel_ptbs))
    pt = ptb.PathTracer(0)
    sc.upload_scene(pt, scene)
    pt.build_bvh()
    wall, dev = pt.build_subgrid(8, 8, 50, gpu=True)
    print(f"Lucy stand-in Subgrid 8/8/50: {pt.structure_bytes(MODE_TWO_PLANE) / 1e9:.2f} GB, GPU build {wall / 1e3:.2f} s wall / {dev / 1e3:.2f} s device")
    pt.set_params(width=1920, height=1080, mode=MODE_BVH)
    t1, tri1, _ = pt.primary_hits()
    pt.set_params(mode=MODE_TWO_PLANE)
    t2, tri2, _ = pt.primary_hits()
    pt.close()
    assert int((_bits(t1) != _bits(t2)).sum()) == 0
    assert float((tri1 == tri2).mean()) >= 0.9999
