"""The BVH built on the GPU (ptb_build_bvh_gpu, csrc/ptb_lbvh.cu) against the host SAH tree, the
oracle and the reference's golden hits.  Closest hits do not depend on the tree, so the bar is the same as for
every other structure: hit distances bit-equal, ids equal (equal distances are resolved by triangle id on the GPU,
so two trees may not even differ on ties), frames / RNG streams / sample counts bit-identical."""
import hashlib
import os

import numpy as np
import pytest

from oracle_lib import MODE_ALL, MODE_BVH, MODE_SUB_BEAMS, MODE_TWO_PLANE, Oracle

pytestmark = pytest.mark.gpu


def _bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def _same_hits(t, tri, rt, rtri, what):
    bad = int((_bits(t) != _bits(rt)).sum())
    assert bad == 0, f"{what}: {bad} rays with a different hit distance"
    assert int((tri != rtri).sum()) == 0, f"{what}: {int((tri != rtri).sum())} rays with a different triangle id"


def test_cornell_gpu_tree_matches_oracle_and_sah_frames(ptb, cornel_ptbs):
    W = H = 256
    o = Oracle(cornel_ptbs)
    o.build()
    o.configure(W, H, MODE_BVH, nee=1, max_depth=0)
    rt, rtri, _ = o.primary_hits()
    rimg, rs, rrng, _ = o.render(3)
    o.close()
    frames = []
    for gpu in (False, True):
        pt = ptb.PathTracer(0)
        pt.load_ptbs(cornel_ptbs)
        ms = pt.build_bvh(gpu=gpu)
        assert ms > 0.0 and 0 < pt.bvh_depth() < 64
        assert pt.structure_bytes(MODE_BVH) > 202 * 48
        pt.set_params(width=W, height=H, mode=MODE_BVH, nee=1, max_depth=0)
        t, tri, _ = pt.primary_hits()
        assert int((_bits(t) != _bits(rt)).sum()) == 0
        assert float((tri == rtri).mean()) >= 0.9999
        pt.reset()
        pt.render(3)
        s, rng = pt.read_state()
        frames.append((pt.read_image(), s, rng, t, tri))
        pt.close()
    (a, sa, ra, ta, tria), (b, sb, rb, tb, trib) = frames
    _same_hits(tb, trib, ta, tria, "GPU tree vs SAH tree")
    assert (_bits(a) == _bits(b)).all() and (sa == sb).all() and (ra == rb).all(), "frames differ between the two trees"
    assert int((rb != rrng).sum()) <= W * H // 2000 and (sb == rs).all()
    assert int((_bits(b) != _bits(rimg)).any(axis=-1).sum()) <= W * H // 2000


def test_gpu_tree_feeds_subgrid_and_subbeams(ptb, cornel_ptbs):
    """The Subgrid hierarchy derives from the BVH (lightFieldGrid.cpp:189-322): built over the device tree it must
    still return the same closest hits; Subbeams only need the leaf order and the scene box."""
    W = H = 192
    hits = {}
    for gpu in (False, True):
        pt = ptb.PathTracer(0)
        pt.load_ptbs(cornel_ptbs)
        pt.build_bvh(gpu=gpu)
        pt.build_subgrid(10, 5, 40)
        pt.build_subbeams(10, 40)
        for mode in (MODE_ALL, MODE_BVH, MODE_TWO_PLANE, MODE_SUB_BEAMS):
            pt.set_params(width=W, height=H, mode=mode)
            t, tri, _ = pt.primary_hits()
            hits[(gpu, mode)] = (t, tri)
        pt.close()
    ref_t, ref_tri = hits[(False, MODE_ALL)]
    for key, (t, tri) in hits.items():
        assert int((_bits(t) != _bits(ref_t)).sum()) == 0, f"{key}: distances differ from brute force"
        assert float((tri == ref_tri).mean()) >= 0.9999, key


@pytest.mark.parametrize("n", [0, 1, 2, 3, 4, 7, 64])
def test_tiny_and_degenerate_scenes(ptb, n):
    """Empty scene, a single leaf (n <= 3), the first real inner node (n = 4) and n identical triangles (equal Morton
    keys: the hierarchy falls back to the sorted position, depth log2 n)."""
    sc = ptb.scenes
    rs = np.random.RandomState(n)
    pos = np.zeros((n, 9), np.float32)
    if n:
        pos[:] = np.array([-1, -1, 5, 1, -1, 5, 0, 1, 5], np.float32)
        if n == 7:
            pos += rs.uniform(-0.5, 0.5, size=(n, 1)).astype(np.float32)  # shifted copies, distinct keys
    scene = sc.scene_from_triangles(pos)
    rays = np.zeros((33, 6), np.float32)
    rays[:, 0] = np.linspace(-1.5, 1.5, 33)
    rays[:, 5] = 1.0
    res = []
    for gpu in (False, True):
        pt = ptb.PathTracer(0)
        sc.upload_scene(pt, scene)
        pt.build_bvh(gpu=gpu)
        res.append(pt.trace_rays(rays, MODE_BVH))
        if gpu and n:
            ta, tria = pt.trace_rays(rays, MODE_ALL)
            _same_hits(res[-1][0], res[-1][1], ta, tria, f"n={n} GPU tree vs brute force")
        pt.close()
    _same_hits(res[1][0], res[1][1], res[0][0], res[0][1], f"n={n}")
    if n == 0:
        assert np.isinf(res[1][0]).all() and (res[1][1] == -1).all()
    else:
        assert (res[1][0] < np.inf).sum() >= 10


def test_soup_gpu_tree_vs_sah_and_oracle(ptb, tmp_path):
    sc = ptb.scenes
    scene = sc.triangle_soup(100_000)
    rays = sc.soup_rays(1 << 16)
    path = str(tmp_path / "soup.ptbs")
    sc.write_ptbs(path, sc.triangle_soup(10_000))
    o = Oracle(path)
    o.build()
    o.configure(64, 64, MODE_BVH)
    rt, rtri = o.trace_rays(rays[:1 << 13])
    o.close()
    pt = ptb.PathTracer(0)
    sc.upload_scene(pt, sc.triangle_soup(10_000))
    pt.build_bvh(gpu=True)
    t, tri = pt.trace_rays(rays[:1 << 13], MODE_BVH)
    assert int((_bits(t) != _bits(rt)).sum()) == 0 and float((tri == rtri).mean()) >= 0.9999
    res = []
    for gpu in (False, True):
        sc.upload_scene(pt, scene)
        ms = pt.build_bvh(gpu=gpu)
        res.append(pt.trace_rays(rays, MODE_BVH) + (ms, pt.bvh_depth()))
    pt.close()
    _same_hits(res[1][0], res[1][1], res[0][0], res[0][1], "soup 1e5")
    print(f"soup 1e5: host SAH {res[0][2]:.1f} ms depth {res[0][3]}, GPU {res[1][2]:.2f} ms depth {res[1][3]}")


def test_lucy_full_size_gpu_tree_against_reference_golden(ptb, cornel_ptbs):
    """BASELINE config 3 at full size over the device-built tree: the reference's own hit distances (digest) and ids."""
    gold = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "lucy_standin_ref_1080p.npz"))
    sc = ptb.scenes
    scene = sc.lucy_standin(sc.read_ptbs(cornel_ptbs))
    pt = ptb.PathTracer(0)
    sc.upload_scene(pt, scene)
    ms = pt.build_bvh(gpu=True)
    pt.set_params(width=1920, height=1080, mode=MODE_BVH)
    t, tri, _ = pt.primary_hits()
    # a full path over the device tree = the same frame as over the SAH tree
    pt.set_params(width=480, height=272, mode=MODE_BVH, nee=1, max_depth=9)
    pt.reset()
    pt.render(2)
    a, (sa, ra) = pt.read_image(), pt.read_state()
    depth = pt.bvh_depth()
    pt.build_bvh(gpu=False)
    pt.reset()
    pt.render(2)
    b, (sb, rb) = pt.read_image(), pt.read_state()
    pt.close()
    print(f"Lucy stand-in: GPU build {ms:.2f} ms, depth {depth}")
    assert hashlib.sha256(t.tobytes()).hexdigest() == str(gold["t_sha256"]), "hit distances differ from the reference's"
    assert float((tri == gold["tri"]).mean()) >= 0.9999
    assert (_bits(a) == _bits(b)).all() and (sa == sb).all() and (ra == rb).all()
