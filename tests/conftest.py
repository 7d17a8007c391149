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
