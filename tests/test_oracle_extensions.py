"""The oracle's preconditioned / warm-started loop (oracle_pcg) is not in the reference; it is
pinned by construction: without a preconditioner and without an initial guess it must give the
bits of oracle_conj_grad (which is pinned to the compiled reference), and with them it must agree
with CG on the equivalent transformed systems."""
import numpy as np
import pytest

import helpers as H

SYSTEMS = [lambda: H.lap(12, 9), lambda: H.lap(6, 5, 4), lambda: H.banded(301, 5), lambda: H.dense(48),
           lambda: H.tridiag4()]


@pytest.mark.parametrize("mk", SYSTEMS)
@pytest.mark.parametrize("iters,tol", [(0, 0.0), (7, 0.0), (500, 1e-10)])
def test_plain_pcg_is_bitwise_the_pinned_loop(mk, iters, tol):
    s = mk()
    if s.name == "tridiag4" and iters > 1 and tol == 0.0:
        iters = 1
    x0, it0, rr0, st0 = H.oracle_cg(s, iters, tol=tol, guard=1)
    x1, it1, rr1, st1 = H.oracle_pcg(s, iters, tol=tol)
    assert it0 == it1 and st0 == st1
    assert np.array_equal(x0.view(np.uint64), x1.view(np.uint64))
    assert rr0 == rr1


def test_warm_start_equals_correction_solve():
    """CG from x0 solves A d = b - A x0 and returns x0 + d: compare with that run of the pinned loop."""
    s = H.banded(257, 4)
    rng = np.random.default_rng(5)
    x0 = rng.standard_normal(s.n)
    x, it, rr, st = H.oracle_pcg(s, 300, tol=1e-12, x0=x0)
    Ax0 = np.empty(s.n)
    H.oracle().oracle_spmv(s.n, H.iptr(s.row_ptr), H.iptr(s.cols), H.dptr(s.vals), H.dptr(x0), H.dptr(Ax0))
    corr = H.System(s.row_ptr, s.cols, s.vals, s.b - Ax0)
    d, itd, _, _ = H.oracle_cg(corr, 300, tol=1e-12 * np.linalg.norm(s.b) / np.linalg.norm(corr.b), guard=1)
    assert abs(it - itd) <= 1
    assert np.linalg.norm(x - (x0 + d)) <= 1e-10 * np.linalg.norm(x)
    assert H.true_relres(s, x) <= 2e-12


def test_warm_start_at_the_solution_stops_at_once():
    s = H.lap(10, 10)
    xs, _, _, _ = H.oracle_cg(s, 1000, tol=1e-14, guard=1)
    x, it, rr, st = H.oracle_pcg(s, 50, tol=1e-10, x0=xs)
    assert it == 0 and st == 1 and np.array_equal(x, xs)


def test_jacobi_equals_cg_on_the_scaled_system():
    """Jacobi-PCG on A is CG on D^-1/2 A D^-1/2 (x = D^-1/2 y): same iterates up to rounding."""
    n = 200
    rng = np.random.default_rng(11)
    scale = 10.0 ** rng.uniform(-2, 2, n)          # badly scaled rows/columns
    base = H.banded(n, 3)
    rp, ci = base.row_ptr, base.cols
    rows = np.repeat(np.arange(n), np.diff(rp))
    vals = base.vals * scale[rows] * scale[ci]
    s = H.System(rp, ci, vals, base.b * scale)
    x, it, rr, st = H.oracle_pcg(s, 25, jacobi=True)
    dh = np.sqrt(H.diag_inv(s))
    t = H.System(rp, ci, vals * dh[rows] * dh[ci], s.b * dh)
    y, ity, _, _ = H.oracle_cg(t, 25, guard=1)
    assert it == ity
    assert np.linalg.norm(x - dh * y) <= 1e-9 * np.linalg.norm(x)
    # and it pays: the unpreconditioned loop needs far more iterations on this system
    _, it_plain, _, _ = H.oracle_cg(s, 5000, tol=1e-10, guard=1)
    _, it_jac, _, stj = H.oracle_pcg(s, 5000, tol=1e-10, jacobi=True)
    assert stj == 1 and it_jac * 3 < it_plain
