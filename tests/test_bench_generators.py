"""CPU: the host-side generator bench.py uses for its long-row line (banded SPD, numpy only)."""
import os
import sys

import numpy as np

import helpers as H

sys.path.insert(0, H.ROOT)
import bench  # noqa: E402


def test_banded_host_system_is_a_valid_spd_csr_system():
    n, hb = 3000, 40
    rp, cols, vals, b = bench.banded_host_system(n, hb)
    assert rp.dtype == np.int32 and cols.dtype == np.int32 and vals.dtype == np.float64
    assert rp[0] == 0 and rp[-1] == len(vals) == len(cols) and len(b) == n
    lens = np.diff(rp)
    assert lens.max() == 2 * hb + 1 and lens[0] == hb + 1 and lens[-1] == hb + 1
    # ascending unique 0-based columns per row: the reference's precondition (mv_ops.c:106-110)
    for i in (0, 1, hb, n // 2, n - 1):
        c = cols[rp[i]:rp[i + 1]]
        assert np.all(np.diff(c) > 0) and c[0] == max(0, i - hb) and c[-1] == min(n - 1, i + hb)
    A = np.zeros((n, n))
    for i in range(n):
        A[i, cols[rp[i]:rp[i + 1]]] = vals[rp[i]:rp[i + 1]]
    assert np.array_equal(A, A.T)
    off = np.abs(A).sum(axis=1) - np.abs(np.diag(A))
    assert np.all(np.diag(A) > off)                       # strictly diagonally dominant => SPD
    # and it is the kind of system the oracle solves in a few iterations
    s = H.System(rp, cols, vals, b)
    x, it, rr, _ = H.oracle_cg(s, 200, 1e-10)
    assert it < 40 and np.linalg.norm(A @ x - b) <= 1e-9 * np.linalg.norm(b)


def test_banded_host_system_is_deterministic():
    a = bench.banded_host_system(500, 7)
    b = bench.banded_host_system(500, 7)
    assert all(np.array_equal(u, v) for u, v in zip(a, b))


def test_both_arms_name_the_workload_identically():
    # the driver compares the `config` dict of the b200 arm with the reference arm's: nothing in it
    # may depend on how a run was partitioned (that lives under `run`)
    for nm in ("dense32768", "lap2d4096", "lap3d512", "lap3d384w"):
        assert bench.workload_config(nm) == bench.workload_config(nm, 8)
        assert "partition" not in bench.workload_config(nm)

