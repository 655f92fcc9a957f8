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


@pytest.fixture(autouse=True)
def _guard_bands(request):
    """Under CGB200_GUARD=1 (tests/test_gpu_guard.py re-runs the GPU suites that way) every device
    buffer of the library sits between poisoned guard bands: a test that made a kernel store out
    of bounds fails here, by name."""
    yield
    if os.environ.get("CGB200_GUARD", "0") in ("", "0") or "gpu" not in request.keywords:
        return
    import __graft_entry__ as g

    pkg_ = g.load_package()
    bad = pkg_.lib().cgb200_guard_check()
    assert bad == 0, f"guard bands damaged in {bad} device buffer(s): {pkg_.last_error()}"
