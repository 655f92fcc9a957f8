"""GPU parity tests of the solve options beyond the reference: an initial guess
(CGB200_OPT_X0) and the Jacobi preconditioner (CGB200_OPT_JACOBI), against oracle_pcg
(oracle/cg_oracle.c), which is itself pinned bitwise to the reference-pinned loop when both are off
(tests/test_oracle_extensions.py).  Same bars as the main parity tests."""
import ctypes as C

import numpy as np
import pytest

import helpers as H

pytestmark = pytest.mark.gpu

TOL_X = 1e-8
TOL_RES = 1e-10
ITER_SLACK = 2


def _dev(pkg, s):
    return pkg.DeviceSystem.from_csr(s.n, 0, s.row_ptr, s.cols, s.vals, s.b)


def _rel(a, b):
    return np.linalg.norm(a - b) / np.linalg.norm(b)


def _badly_scaled(n=3000, hb=4, seed=11):
    """Banded SPD system whose rows/columns are scaled over four decades: Jacobi undoes the scaling."""
    rng = np.random.default_rng(seed)
    scale = 10.0 ** rng.uniform(-2, 2, n)
    base = H.banded(n, hb)
    rows = np.repeat(np.arange(n), np.diff(base.row_ptr))
    return H.System(base.row_ptr, base.cols, base.vals * scale[rows] * scale[base.cols], base.b * scale,
                    name=f"scaled_banded_{n}")


SYSTEMS = [
    lambda: H.lap(40, 33),
    lambda: H.lap(17, 16, 15),
    lambda: H.banded(5001, 6),
    lambda: H.dense(700),
    lambda: _badly_scaled(),
]


@pytest.mark.parametrize("mk", SYSTEMS)
@pytest.mark.parametrize("jacobi", [False, True])
@pytest.mark.parametrize("warm", [False, True])
def test_fixed_count_matches_oracle(pkg, gpu_lib, mk, jacobi, warm):
    s = mk()
    x0 = np.random.default_rng(3).standard_normal(s.n) if warm else None
    d = _dev(pkg, s)
    try:
        for iters in (0, 9):
            xo, it_o, rr_o, _ = H.oracle_pcg(s, iters, x0=x0, jacobi=jacobi)
            r = d.solve(iters, jacobi=jacobi, x0=x0)
            assert r.iterations == it_o and r.status == pkg.MAXITER
            assert _rel(d.x(), xo) <= TOL_X
            assert abs(r.rr - rr_o) <= 1e-6 * rr_o        # the reported r.r is the true recursive one
    finally:
        d.free()


@pytest.mark.parametrize("mk", SYSTEMS)
def test_options_off_is_the_reference_loop(pkg, gpu_lib, mk):
    # flags = 0 must not change anything: same bits as the default path, run to run
    s = mk()
    d = _dev(pkg, s)
    try:
        d.solve(7)
        x_ref = d.x()
        d.solve(3, jacobi=True, x0=x_ref)    # run the other path in between
        d.solve(7)
        assert np.array_equal(x_ref.view(np.uint64), d.x().view(np.uint64))
    finally:
        d.free()


@pytest.mark.parametrize("mk", SYSTEMS)
@pytest.mark.parametrize("jacobi", [False, True])
def test_tolerance_stop(pkg, gpu_lib, mk, jacobi):
    s = mk()
    if "scaled" in s.name and not jacobi:
        pytest.skip("80 000 iterations without the preconditioner: the count is not comparable to +-2")
    xo, it_o, _, st_o = H.oracle_pcg(s, 20000, tol=TOL_RES, jacobi=jacobi)
    assert st_o == 1
    d = _dev(pkg, s)
    try:
        r = d.solve(20000, tol=TOL_RES, jacobi=jacobi)
        assert r.status == pkg.CONVERGED
        assert abs(r.iterations - it_o) <= ITER_SLACK
        assert d.true_relres() <= 1.05 * TOL_RES
        assert _rel(d.x(), xo) <= TOL_X
    finally:
        d.free()


def test_jacobi_pays_on_a_badly_scaled_system(pkg, gpu_lib):
    s = _badly_scaled()
    d = _dev(pkg, s)
    try:
        plain = d.solve(100000, tol=TOL_RES)
        jac = d.solve(100000, tol=TOL_RES, jacobi=True)
        assert jac.status == pkg.CONVERGED and d.true_relres() <= 1.05 * TOL_RES
        assert jac.iterations * 3 < plain.iterations, (jac.iterations, plain.iterations)
    finally:
        d.free()


def test_warm_start_from_previous_solution(pkg, gpu_lib):
    # restart: a second solve that starts from the first one's x continues the descent
    s = H.lap(64, 64)
    d = _dev(pkg, s)
    try:
        d.solve(40)
        t1 = d.true_relres()
        d.solve(40, x0=True)
        t2 = d.true_relres()
        assert t2 < 0.5 * t1
        full = d.solve(100000, tol=TOL_RES, x0=True)
        assert full.status == pkg.CONVERGED and d.true_relres() <= 1.05 * TOL_RES
        # starting at the solution: (next to) nothing left to do -- the true residual of the converged
        # x may sit a hair above the recursive one that stopped the first solve
        again = d.solve(100000, tol=TOL_RES, x0=True)
        assert again.iterations <= 2 and again.status == pkg.CONVERGED
    finally:
        d.free()


def test_warm_start_at_exact_solution_is_status_exact(pkg, gpu_lib):
    s = H.tridiag4()
    d = _dev(pkg, s)
    try:
        r = d.solve(10, x0=np.array([2.0, 3.0, 3.0, 2.0]))
        assert r.status == pkg.EXACT and r.iterations == 0
        assert np.array_equal(d.x(), [2.0, 3.0, 3.0, 2.0])
    finally:
        d.free()


def test_device_generated_systems(pkg, gpu_lib):
    # Laplacians have a constant diagonal: Jacobi only rescales, the iterates are CG's
    d = pkg.DeviceSystem.lap(96, 96, 96)
    try:
        d.solve(30)
        x_cg = d.x()
        r = d.solve(30, jacobi=True)
        assert r.iterations == 31
        assert _rel(d.x(), x_cg) <= 1e-9
    finally:
        d.free()
    n = 3000
    s = H.dense(n)
    d = pkg.DeviceSystem.dense(n)
    try:
        xo, it_o, _, _ = H.oracle_pcg(s, 3, jacobi=True)
        r = d.solve(3, jacobi=True)
        assert r.iterations == it_o and _rel(d.x(), xo) <= TOL_X
        assert d.one_kernel_iteration()      # the default dense path is untouched by the option
    finally:
        d.free()


def test_pattern_format_with_jacobi(pkg, gpu_lib):
    s = H.lap(48, 40, 30)
    d = _dev(pkg, s)
    try:
        assert d.use_pattern_format(True)
        xo, it_o, _, _ = H.oracle_pcg(s, 12, jacobi=True)
        r = d.solve(12, jacobi=True)
        assert r.iterations == it_o and _rel(d.x(), xo) <= TOL_X
    finally:
        d.free()


def test_refusals(pkg, gpu_lib):
    s = H.lap(12, 12)
    d = _dev(pkg, s)
    try:
        with pytest.raises(Exception):
            d.solve(5, variant=pkg.SINGLE_REDUCTION, jacobi=True)
        assert "classic" in pkg.last_error()
        o = pkg.Opts(5, 0.0, 0, 0, pkg.OPT_JACOBI)
        t = C.c_longlong(-1)
        assert gpu_lib.cgb200_system_solve_async(d.h, C.byref(o), C.byref(t)) == -1
    finally:
        d.free()
    # a matrix with a non-positive diagonal entry cannot take the Jacobi preconditioner
    bad = H.lap(12, 12)
    vals = bad.vals.copy()
    diag_pos = [k for i in range(bad.n) for k in range(bad.row_ptr[i], bad.row_ptr[i + 1]) if bad.cols[k] == i]
    vals[diag_pos[5]] = -4.0
    d = _dev(pkg, H.System(bad.row_ptr, bad.cols, vals, bad.b))
    try:
        with pytest.raises(Exception):
            d.solve(5, jacobi=True)
        assert "diagonal" in pkg.last_error()
    finally:
        d.free()


def test_struct_api_with_initial_guess(pkg, gpu_lib):
    # cgb200_solve on the reference's structs: *x on entry is the guess
    s = H.banded(2001, 5)
    x0 = np.random.default_rng(9).standard_normal(s.n)
    xo, it_o, _, _ = H.oracle_pcg(s, 11, x0=x0, jacobi=True)
    A = pkg.MvSparse(s.n, s.nnz, H.dptr(s.vals), H.iptr(s.cols), H.iptr(s.row_ptr))
    b = pkg.MvSparse(s.n, s.n, H.dptr(s.b), None, None)
    guess = pkg.MvSparse(s.n, s.n, H.dptr(x0), None, None)
    xp = C.pointer(guess)
    o = pkg.Opts(11, 0.0, 0, 0, pkg.OPT_JACOBI | pkg.OPT_X0)
    r = pkg.Result()
    assert gpu_lib.cgb200_solve(C.byref(A), C.byref(b), C.byref(xp), C.byref(o), C.byref(r)) == 0, pkg.last_error()
    x = np.ctypeslib.as_array(xp.contents.values, shape=(s.n,)).copy()
    gpu_lib.cgb200_invalidate(C.byref(A))
    assert r.iterations == it_o and _rel(x, xo) <= TOL_X
    assert C.addressof(xp.contents) != C.addressof(guess)       # a fresh vector, as conj_grad returns
    # missing guess: refused
    null = pkg.MvSparseP()
    assert gpu_lib.cgb200_solve(C.byref(A), C.byref(b), C.byref(null), C.byref(o), C.byref(r)) == -1
    gpu_lib.cgb200_invalidate(C.byref(A))


# ---------------------------------------------------------------------------------------------
# diagonal (DIA) storage for banded matrices with arbitrary coefficients (csrc/banded.cu)
# ---------------------------------------------------------------------------------------------


def _holey_band(n, hb, drop, seed=11):
    """Symmetric banded SPD system with a fraction `drop` of its off-diagonal pairs missing (the band
    stays mostly full, rows have different lengths, the chain i -- i+1 is kept: irreducible)."""
    rng = np.random.default_rng(seed)
    s = H.banded(n, hb)
    keep = np.ones(s.nnz, dtype=bool)
    rows = np.repeat(np.arange(n), np.diff(s.row_ptr))
    off = s.cols - rows
    pair_lo, pair_hi = np.minimum(rows, s.cols), np.maximum(rows, s.cols)
    h = (pair_lo.astype(np.int64) * 1000003 + pair_hi) % 1000
    keep &= ~((np.abs(off) > 1) & (h < int(drop * 1000)))
    rp = np.zeros(n + 1, dtype=np.int64)
    np.add.at(rp, rows[keep] + 1, 1)
    return H.System(np.cumsum(rp), s.cols[keep], s.vals[keep], s.b, "holey_band")


BANDED_SYSTEMS = [lambda: H.banded(3000, 40), lambda: H.banded(700, 300), lambda: H.banded(4097, 5),
                  lambda: _holey_band(2500, 60, 0.2), lambda: H.lap(1, 2000)]


@pytest.mark.parametrize("mk", BANDED_SYSTEMS)
def test_banded_format_matvec_bit_exact_and_solves(pkg, gpu_lib, mk):
    s = mk()
    d = pkg.DeviceSystem.from_csr(s.n, 0, s.row_ptr, s.cols, s.vals, s.b)
    it_b0, k1_b0 = d.bytes()
    assert d.use_banded_format(True), pkg.last_error()
    it_b1, _ = d.bytes()
    assert it_b1 < it_b0                                     # fewer bytes per iteration than CSR
    want = np.empty(s.n)
    for seed in (3, 4):
        v = H.rhs(s.n, seed)
        H.oracle().oracle_spmv(s.n, H.iptr(s.row_ptr), H.iptr(s.cols), H.dptr(s.vals), H.dptr(v), H.dptr(want))
        assert np.array_equal(d.matvec(v).view(np.uint64), want.view(np.uint64))
    K = 12
    xo, it, _, _ = H.oracle_cg(s, K)
    r = d.solve(K)
    assert r.iterations == it
    x_dia = d.x()
    assert np.linalg.norm(x_dia - xo) <= 1e-8 * np.linalg.norm(xo)
    rt = d.solve(5000, tol=1e-10)
    assert rt.status == pkg.CONVERGED and d.true_relres() <= 1.05e-10
    rj = d.solve(5000, tol=1e-10, jacobi=True)               # the preconditioned loop takes the same matvec
    assert rj.status == pkg.CONVERGED
    assert not d.use_banded_format(False)                    # back on CSR: same x to the last bits of the dots
    d.solve(K)
    assert np.linalg.norm(d.x() - x_dia) <= 1e-12 * np.linalg.norm(x_dia)
    d.free()


def test_banded_format_refuses_what_is_not_a_mostly_full_band(pkg, gpu_lib):
    s = H.lap(64, 64)                                        # 5 entries in a band of 129 diagonals
    d = pkg.DeviceSystem.from_csr(s.n, 0, s.row_ptr, s.cols, s.vals, s.b)
    assert not d.use_banded_format(True)
    assert "not applicable" in pkg.last_error() and "full" in pkg.last_error()
    xo, it, _, _ = H.oracle_cg(s, 30)                        # it stays on CSR and keeps working
    d.solve(30)
    assert np.linalg.norm(d.x() - xo) <= 1e-8 * np.linalg.norm(xo)
    d.free()
    dd = pkg.DeviceSystem.dense(256)                         # the dense path has no use for it
    assert not dd.use_banded_format(True)
    dd.free()

