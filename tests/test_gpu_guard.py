"""Memory-safety pass over the GPU suites in the library's guard mode.

compute-sanitizer is not available on the GPU pool this repository is developed on, so the library
carries its own check (csrc/devmem.h): with CGB200_GUARD=1 every device buffer it allocates sits
between two 4 KiB guard bands and is filled -- guards and payload -- with 0xFF.  A kernel that
stores out of bounds damages a guard (tests/conftest.py checks after every GPU test and names the
allocation site); one that loads out of bounds or loads memory nobody wrote gets NaN doubles / -1
indices, which the parity comparisons against the oracle cannot miss.  The variable is read when
the library makes its first allocation, hence the fresh interpreter.
"""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SUITES = ["tests/test_gpu_parity.py", "tests/test_gpu_dropin.py", "tests/test_gpu_extensions.py"]


def _run(args, **env):
    e = dict(os.environ, CGB200_GUARD="1", **env)
    return subprocess.run([sys.executable, *args], cwd=ROOT, env=e, stdout=subprocess.PIPE,
                          stderr=subprocess.STDOUT, text=True, timeout=1500)


@pytest.mark.gpu
def test_guard_mode_catches_a_store_outside_a_buffer(gpu_lib):
    code = (
        "import __graft_entry__ as g; p = g.load_package(); L = p.lib(); "
        "assert L.cgb200_init(-1) == 0, p.last_error(); "
        "assert L.cgb200_guard_check() == 0; "
        "r = L.cgb200_guard_selftest(); assert r == 1, r; "
        "assert L.cgb200_guard_check() == 0; print('guard selftest ok')"
    )
    out = _run(["-c", code])
    assert out.returncode == 0 and "guard selftest ok" in out.stdout, out.stdout[-2000:]


@pytest.mark.gpu
def test_guard_mode_is_off_by_default(gpu_lib):
    if os.environ.get("CGB200_GUARD", "0") not in ("", "0"):
        pytest.skip("this session itself runs in guard mode")
    assert gpu_lib.cgb200_guard_check() == -1
    assert gpu_lib.cgb200_guard_selftest() == -1


@pytest.mark.gpu
def test_gpu_suites_pass_with_guard_bands_and_poisoned_buffers(gpu_lib):
    if os.environ.get("CGB200_GUARD", "0") not in ("", "0"):
        pytest.skip("already inside the guarded run")
    out = _run(["-m", "pytest", *SUITES, "-q", "-x", "-m", "gpu", "-p", "no:cacheprovider"])
    tail = out.stdout[-3000:]
    assert out.returncode == 0, tail
    assert " passed" in tail and "failed" not in tail, tail


@pytest.mark.gpu
def test_stream_kernel_keeps_its_long_row_coverage(gpu_lib):
    # single-GPU blocks with a mean row length >= 64 take the row-block kernel (16 rows per warp);
    # multi-rank systems still run such rows through the stream kernel's chunked-row path, so the
    # same bit-exactness tests are repeated with the selection switched off -- and with the other
    # selectable long-row forms (the shuffle-chain kernel, 32- and 8-row units)
    for extra in ({"CGB200_SPMV_ROWGROUP_MIN": "0"},
                  {"CGB200_SPMV_LONGROW": "group", "CGB200_SPMV_ROWGROUP_MIN": "16"},
                  {"CGB200_SPMV_LONGROW": "block", "CGB200_SPMV_ROWGROUP_MIN": "16"},
                  {"CGB200_SPMV_LONGROW": "block8", "CGB200_SPMV_ROWGROUP_MIN": "16"}):
        _long_row_tests(extra)


def _long_row_tests(extra):
    e = dict(os.environ, **extra)
    out = subprocess.run([sys.executable, "-m", "pytest", "tests/test_gpu_parity.py", "-q", "-x", "-m", "gpu",
                          "-p", "no:cacheprovider", "-k", "csr_bit_exact or ragged or random_sparse or long_row"],
                         cwd=ROOT, env=e, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900)
    tail = out.stdout[-3000:]
    assert out.returncode == 0 and " passed" in tail, tail
