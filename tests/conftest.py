import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def pkg():
    import __graft_entry__ as g

    return g.load_package()


@pytest.fixture(scope="session")
def gpu_lib(pkg):
    """libcgb200.so on a box with a GPU; fails loudly (no skip, no fallback) when there is none."""
    L = pkg.lib()
    assert L.cgb200_device_available() == 1, "gpu test on a box without a CUDA device"
    assert L.cgb200_init(-1) == 0, pkg.last_error()
    return L
