"""The Subbeam structure built level by level on the GPU (ptb_build_subbeams_gpu, csrc/ptb_beam_gpu.cu over
csrc/ptb_beam_levels.h) against the depth-first host builder: the same functors were checked against
build_subbeams with plain loops (tests/test_native_host.py), here they run as kernels — the bar is byte equality of beam
records, leaf slices, root tables and planes (ptb_structure_digest), plus closest-hit parity of what comes out."""
import numpy as np
import pytest

from oracle_lib import MODE_BVH, MODE_SUB_BEAMS, Oracle

pytestmark = pytest.mark.gpu


def _bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def _digests(pt, gs, m, depth):
    pt.build_subbeams(gs, m, max_depth=depth, gpu=False)
    host = (pt.structure_digest(MODE_SUB_BEAMS), pt.structure_bytes(MODE_SUB_BEAMS))
    wall, dev = pt.build_subbeams(gs, m, max_depth=depth, gpu=True)
    gpu = (pt.structure_digest(MODE_SUB_BEAMS), pt.structure_bytes(MODE_SUB_BEAMS))
    assert host[0] != 0 and gpu[0] != 0
    return host, gpu, wall, dev


@pytest.mark.parametrize("gs,m,depth", [(10, 40, 12), (4, 8, 3), (3, 500, 12)])
def test_cornell_pools_identical_to_host_builder(ptb, cornel_ptbs, gs, m, depth):
    """m = 500 > 202 triangles: every root beam is a leaf (no level below 0); gs 4 / m 8: three levels of splits."""
    pt = ptb.PathTracer(0)
    pt.load_ptbs(cornel_ptbs)
    pt.build_bvh()
    host, gpu, wall, dev = _digests(pt, gs, m, depth)
    print(f"Cornell gs {gs} m {m}: {gpu[1] / 1e6:.2f} MB, GPU build {wall:.1f} ms wall / {dev:.1f} ms stream")
    assert host == gpu, f"pools differ: host {host}, gpu {gpu}"
    if depth == 12 and gs == 10:  # and the structure renders the oracle's frame
        W = H = 96
        o = Oracle(cornel_ptbs)
        o.build(beams=True, bgs=gs, bm=m, max_beam_depth=depth)
        o.configure(W, H, MODE_SUB_BEAMS, nee=1, max_depth=0)
        rimg, rs, rrng, _ = o.render(2)
        o.close()
        pt.set_params(width=W, height=H, mode=MODE_SUB_BEAMS, nee=1, max_depth=0)
        pt.reset()
        pt.render(2)
        s, rng = pt.read_state()
        img = pt.read_image()
        assert (s == rs).all() and int((rng != rrng).sum()) <= W * H // 2000
        assert int((_bits(img) != _bits(rimg)).any(axis=-1).sum()) <= W * H // 2000
    pt.close()


def test_soup_empty_scene_and_device_tree(ptb, cornel_ptbs):
    sc = ptb.scenes
    pt = ptb.PathTracer(0)
    sc.upload_scene(pt, sc.triangle_soup(10_000))
    pt.build_bvh(gpu=True)  # leaf order from the device BVH builder
    host, gpu, wall, dev = _digests(pt, 4, 40, 2)
    print(f"soup 1e4 gs 4 m 40 depth 2: {gpu[1] / 1e6:.1f} MB, GPU build {wall:.1f} ms wall / {dev:.1f} ms stream")
    assert host == gpu, "soup: pools differ"
    pt.set_params(width=128, height=128, mode=MODE_BVH)
    t1, tri1, _ = pt.primary_hits()
    pt.set_params(mode=MODE_SUB_BEAMS)
    t2, tri2, _ = pt.primary_hits()
    pt.close()
    assert int((_bits(t1) != _bits(t2)).sum()) == 0 and float((tri1 == tri2).mean()) >= 0.9999
    pt = ptb.PathTracer(0)
    sc.upload_scene(pt, sc.scene_from_triangles(np.zeros((0, 9), np.float32)))
    pt.build_bvh()
    host, gpu, _, _ = _digests(pt, 3, 40, 12)
    assert host == gpu
    pt.set_params(width=16, height=8, mode=MODE_SUB_BEAMS)
    t, tri, _ = pt.primary_hits()
    assert np.isinf(t).all()
    pt.close()


def test_one_thread_per_beam_variant_in_its_own_process(cornel_ptbs):
    """The chunked workers below the roots (one thread per beam and 64 entries of the parent's list) are the default since
    round 2 — every other test here builds with them; PTB_BEAM_DEEP_CHUNK=0 selects the older one-thread-per-beam steps, which
    must give the host builder's pools too.  Runs in a child process: the knob is read from the environment."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = (
        "import sys; sys.path.insert(0, %r); import ptb_b200 as p\n"
        "pt = p.PathTracer(0); pt.load_ptbs(%r); pt.build_bvh()\n"
        "pt.build_subbeams(4, 8, max_depth=3, gpu=False); h = pt.structure_digest(3)\n"
        "wall, dev = pt.build_subbeams(4, 8, max_depth=3, gpu=True); g = pt.structure_digest(3)\n"
        "print('DEEP', h == g and h != 0, wall, dev)\n" % (root, cornel_ptbs))
    env = dict(os.environ, PTB_BEAM_DEEP_CHUNK="0")
    out = subprocess.run([sys.executable, "-c", code], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    print(out.stdout[-2000:])
    assert out.returncode == 0 and "DEEP True" in out.stdout
