"""Row a16: the display transform (reference src/window/window.cpp:507-531) is byte work and must be byte-exact.
The checker is tests/native/tonemap_check.cpp — the reference's expression compiled with the reference's flags against the
same libm.  CPU: the step thresholds the library hands its kernel equal the ones the checker finds by walking ALL floats of
[0, 1] (which also proves the step function monotone, the one assumption of the threshold search).  GPU: ptb_read_rgba8
against the checker on a rendered frame and on a frame of edge values, zero differing bytes."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


from conftest import apply_display as _apply


@pytest.fixture(scope="module")
def checker(tonemap_checker):
    return tonemap_checker


def test_thresholds_equal_the_exhaustive_walk(ptb, checker, tmp_path):
    out = str(tmp_path / "steps.bin")
    r = subprocess.run([checker, "steps", out], stdout=subprocess.PIPE, text=True)
    assert r.returncode == 0 and "monotone" in r.stdout and "NOT" not in r.stdout, r.stdout
    want = np.fromfile(out, np.float32)
    got = np.zeros(256, np.float32)
    assert ptb.load_library().ptb_tone_thresholds(got.ctypes.data_as(ctypes.c_void_p)) == 0
    assert (got.view(np.uint32) == want.view(np.uint32)).all()
    assert (np.diff(got[1:]) > 0).all() and got[255] <= 1.0


@pytest.mark.gpu
def test_rgba8_is_byte_exact(ptb, cornel_ptbs, checker, tmp_path):
    W = H = 256
    pt = ptb.PathTracer(0)
    pt.load_ptbs(cornel_ptbs)
    pt.build_bvh()
    pt.set_params(width=W, height=H, mode=1, nee=1, max_depth=6)
    pt.reset()
    pt.render(8)
    frames = [pt.read_image().copy()]
    # edge values: for every byte step the radiance values around c = v / (1 - v), a log-uniform cloud, and the specials
    th = np.zeros(256, np.float32)
    pt.L.ptb_tone_thresholds(th.ctypes.data_as(ctypes.c_void_p))
    v = th[1:].astype(np.float64)
    c = (v / np.maximum(1.0 - v, 1e-12)).astype(np.float32)
    near = np.concatenate([np.nextafter(c, np.float32(0)), c, np.nextafter(c, np.float32(np.inf))])
    for _ in range(3):
        near = np.concatenate([near, np.nextafter(near, np.float32(0)), np.nextafter(near, np.float32(np.inf))])
    rng = np.random.default_rng(5)
    cloud = np.exp(rng.uniform(np.log(1e-7), np.log(1e5), W * H * 3)).astype(np.float32)
    special = np.array([0.0, -0.0, 1.0, 0.5, 1e30, 3.4e38, np.inf, np.nan, 1e-45, 1e-38, 254.0 / 255, 255.0 / 255, 1.0000001], np.float32)
    vals = np.concatenate([near, special, cloud])[: W * H * 3]
    edge = np.ones((H, W, 4), np.float32)
    edge[..., :3] = vals.reshape(H, W, 3)
    samples, rng_state = pt.read_state()
    pt.write_state(edge, samples, rng_state)
    frames.append(edge)
    got_edge = {tone: pt.read_rgba8(tone_mapping=bool(tone)).copy() for tone in (1, 0)}
    pt.write_state(frames[0], samples, rng_state)
    got_frame = {tone: pt.read_rgba8(tone_mapping=bool(tone)).copy() for tone in (1, 0)}
    pt.close()
    for name, img, got in (("rendered frame", frames[0], got_frame), ("edge values", edge, got_edge)):
        for tone in (1, 0):
            want = _apply(checker, img, tone, str(tmp_path))
            bad = int((want != got[tone].reshape(H, W, 4)).sum())
            print(f"display transform, {name}, tone mapping {tone}: {bad} of {W * H * 4} bytes differ")
            assert bad == 0
