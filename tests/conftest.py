import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def cornel_ptbs():
    return os.path.join(GOLDEN, "cornel.ptbs")


@pytest.fixture(scope="session")
def ptb():
    import ptb_b200
    ptb_b200.build_native()
    return ptb_b200


@pytest.fixture(scope="session")
def tonemap_checker(tmp_path_factory):
    """tests/native/tonemap_check.cpp: the reference's display transform (window.cpp:507-531), g++ with the reference's flags"""
    import subprocess
    exe = str(tmp_path_factory.mktemp("tone") / "tonemap_check")
    subprocess.check_call(["g++", "-std=gnu++17", "-O3", "-msse", "-msse2", "-mavx", "-mavx2", "-fno-fast-math", "-ffp-contract=off",
                           "-pthread", os.path.join(ROOT, "tests", "native", "tonemap_check.cpp"), "-o", exe])
    return exe


def apply_display(checker, img, tone, workdir):
    """(H, W, 4) float32 frame -> (H, W, 4) bytes as the reference's window would show them"""
    import subprocess
    import numpy as np
    i, o = os.path.join(workdir, "tone_in.bin"), os.path.join(workdir, "tone_out.bin")
    np.ascontiguousarray(img, np.float32).tofile(i)
    subprocess.check_call([checker, "apply", i, o, str(int(tone))])
    return np.fromfile(o, np.uint8).reshape(img.shape[0], img.shape[1], 4)
