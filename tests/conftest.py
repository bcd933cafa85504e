import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "bwa-mem2-arm_b200"), os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def ora():
    import ora as _ora
    _ora.lib()
    return _ora


@pytest.fixture(scope="session")
def bsw():
    """The product binding.  GPU tests fail (never skip to a fallback) when it cannot load."""
    import bswb200
    bswb200.lib()
    return bswb200
