"""GPU parity tests: the CUDA path, called through the C ABI (libcgb200.so), against the CPU oracle.

Bars (BASELINE.json north_star):
  * element-wise operators and the CSR matvec: BIT-EXACT (the kernels keep the reference's
    unfused multiply/add and per-row ascending-column order);
  * dot products: relative error <= 4 n eps (summation order differs);
  * solves: ||x_gpu - x_cpu|| / ||x_cpu|| <= 1e-8 at equal iteration count; with a tolerance,
    iteration count within +-2 of the oracle and true relative residual <= 1e-10 (+5% slack for
    the gap between the recursive and the true residual).
"""
import ctypes as C

import numpy as np
import pytest

import helpers as H

pytestmark = pytest.mark.gpu

EPS = np.finfo(np.float64).eps
TOL_X = 1e-8        # north_star: solution relative L2 difference
TOL_RES = 1e-10     # north_star: relative residual at convergence
ITER_SLACK = 2      # north_star: iteration count within +-2


def _structs(pkg, sysm):
    A = pkg.MvSparse(sysm.n, sysm.nnz, H.dptr(sysm.vals), H.iptr(sysm.cols), H.iptr(sysm.row_ptr))
    b = pkg.MvSparse(sysm.n, sysm.n, H.dptr(sysm.b), None, None)
    return A, b


def _vec(pkg, arr):
    return pkg.MvSparse(len(arr), len(arr), H.dptr(arr), None, None)


def _take(ptr, n):
    return np.ctypeslib.as_array(ptr.contents.values, shape=(n,)).copy() if n else np.empty(0)


def _bits(a):
    return np.ascontiguousarray(a).view(np.uint64)


def gpu_cg(pkg, L, sysm, max_iter):
    A, b = _structs(pkg, sysm)
    xp = pkg.MvSparseP()
    assert L.conj_grad(max_iter, C.byref(A), C.byref(b), C.byref(xp)) == 0, pkg.last_error()
    x = _take(xp, sysm.n)
    L.cgb200_invalidate(C.byref(A))
    return x


def gpu_solve(pkg, L, sysm, max_iter, tol, variant=0):
    A, b = _structs(pkg, sysm)
    xp = pkg.MvSparseP()
    o = pkg.Opts(max_iter, tol, variant, 0, 0)
    r = pkg.Result()
    assert L.cgb200_solve(C.byref(A), C.byref(b), C.byref(xp), C.byref(o), C.byref(r)) == 0, pkg.last_error()
    x = _take(xp, sysm.n)
    L.cgb200_invalidate(C.byref(A))
    return x, r


# ---------------------------------------------------------------------------------------------
# known answers (confirmed on the compiled reference by tests/golden/make_golden.py)
# ---------------------------------------------------------------------------------------------


def test_kat_tridiag4(pkg, gpu_lib):
    s = H.tridiag4()
    assert np.array_equal(gpu_cg(pkg, gpu_lib, s, 0), [2.0, 2.0, 2.0, 2.0])
    assert np.array_equal(gpu_cg(pkg, gpu_lib, s, 1), [2.0, 3.0, 3.0, 2.0])


def test_past_exact_convergence_stops_instead_of_nan(pkg, gpu_lib):
    # the reference returns NaN here (0/0, cg.c:113); the device loop stops with status EXACT
    s = H.tridiag4()
    x, r = gpu_solve(pkg, gpu_lib, s, 10, 0.0)
    assert np.array_equal(x, [2.0, 3.0, 3.0, 2.0])
    assert r.status == pkg.EXACT and r.iterations == 2


# ---------------------------------------------------------------------------------------------
# operators
# ---------------------------------------------------------------------------------------------


@pytest.mark.parametrize("n", [0, 1, 2, 3, 31, 1000, 4097, 1 << 20, (1 << 20) + 1])
def test_elementwise_ops_bit_exact(pkg, gpu_lib, n):
    L, O = gpu_lib, H.oracle()
    a, b = H.rhs(n, 11), H.rhs(n, 12)
    va, vb = _vec(pkg, a), _vec(pkg, b)
    ref = np.empty(n)
    for name, args, orc in [
        ("sv_mult", (C.c_double(-1.7), C.byref(va)), lambda: O.oracle_scale(n, -1.7, H.dptr(a), H.dptr(ref))),
        ("vec_add", (C.byref(va), C.byref(vb)), lambda: O.oracle_add(n, H.dptr(a), H.dptr(b), H.dptr(ref))),
        ("vec_sub", (C.byref(va), C.byref(vb)), lambda: O.oracle_sub(n, H.dptr(a), H.dptr(b), H.dptr(ref))),
    ]:
        out = pkg.MvSparseP()
        assert getattr(L, name)(*args, C.byref(out)) == 0, pkg.last_error()
        assert out.contents.size == n and out.contents.nnz == n
        orc()
        assert np.array_equal(_bits(_take(out, n)), _bits(ref)), name


@pytest.mark.parametrize("n", [1, 2, 33, 1000, 65537, 1 << 22])
def test_dot_product(pkg, gpu_lib, n):
    a, b = H.rhs(n, 21), H.rhs(n, 22)
    va, vb = _vec(pkg, a), _vec(pkg, b)
    got = gpu_lib.dot_product(C.byref(va), C.byref(vb))
    want = H.oracle().oracle_dot(n, H.dptr(a), H.dptr(b))
    scale = float(np.dot(np.abs(a), np.abs(b)))
    assert abs(got - want) <= 4 * n * EPS * scale
    # r.r style (same vector twice)
    got2 = gpu_lib.dot_product(C.byref(va), C.byref(va))
    want2 = H.oracle().oracle_dot(n, H.dptr(a), H.dptr(a))
    assert abs(got2 - want2) <= 4 * n * EPS * want2


def test_dot_is_run_to_run_deterministic(pkg, gpu_lib):
    n = (1 << 21) + 3
    a, b = H.rhs(n, 5), H.rhs(n, 6)
    va, vb = _vec(pkg, a), _vec(pkg, b)
    vals = {gpu_lib.dot_product(C.byref(va), C.byref(vb)) for _ in range(5)}
    assert len(vals) == 1


def test_error_conventions(pkg, gpu_lib):
    L = gpu_lib
    a, b = H.rhs(8, 1), H.rhs(9, 2)
    va, vb = _vec(pkg, a), _vec(pkg, b)
    out = pkg.MvSparseP()
    assert L.vec_add(C.byref(va), C.byref(vb), C.byref(out)) == -1        # size mismatch
    assert L.vec_sub(None, C.byref(vb), C.byref(out)) == -1               # NULL
    assert L.sv_mult(2.0, None, C.byref(out)) == -1
    assert L.dot_product(C.byref(va), C.byref(vb)) == -1.0                # mv_ops.c:125-126
    assert L.dot_product(None, C.byref(vb)) == -1.0
    s = H.lap(4, 4)
    A, _ = _structs(pkg, s)
    assert L.mv_mult(C.byref(A), C.byref(va), C.byref(out)) == -1         # A->size != b->size
    assert not L.mv_deep_copy(None)


def test_in_place_aliasing(pkg, gpu_lib):
    # the reference updates r in place: vec_sub(r, t, &r) (cg.c:123)
    L = gpu_lib
    n = 5001
    r0, t = H.rhs(n, 31), H.rhs(n, 32)
    r = L.new_mv_struct_with_size(n)
    C.memmove(r.contents.values, r0.ctypes.data, 8 * n)
    vt = _vec(pkg, t)
    assert L.vec_sub(r, C.byref(vt), C.byref(r)) == 0
    assert np.array_equal(_bits(_take(r, n)), _bits(r0 - t))


def test_output_reuse_and_resize(pkg, gpu_lib):
    L = gpu_lib
    a = H.rhs(100, 1)
    out = pkg.MvSparseP()
    va = _vec(pkg, a)
    assert L.sv_mult(2.0, C.byref(va), C.byref(out)) == 0 and out.contents.size == 100
    a2 = H.rhs(300, 2)
    va2 = _vec(pkg, a2)
    assert L.sv_mult(2.0, C.byref(va2), C.byref(out)) == 0 and out.contents.size == 300
    assert np.array_equal(_take(out, 300), 2.0 * a2)


SPMV_SYSTEMS = [
    lambda: H.tridiag4(),
    lambda: H.lap(32, 32),
    lambda: H.lap(100, 37),
    lambda: H.lap(12, 12, 12),
    lambda: H.lap(33, 17, 9),
    lambda: H.banded(3000, 5),
    lambda: H.banded(700, 300),       # rows of 601 entries
    lambda: H.banded(2600, 1200),     # rows longer than one 2048-entry stage: chunked rows
    lambda: H.lap(256, 256),
]


@pytest.mark.parametrize("mk", SPMV_SYSTEMS)
def test_mv_mult_csr_bit_exact(pkg, gpu_lib, mk):
    s = mk()
    A, _ = _structs(pkg, s)
    v = H.rhs(s.n, 99)
    vv = _vec(pkg, v)
    out = pkg.MvSparseP()
    assert gpu_lib.mv_mult(C.byref(A), C.byref(vv), C.byref(out)) == 0, pkg.last_error()
    want = np.empty(s.n)
    H.oracle().oracle_spmv(s.n, H.iptr(s.row_ptr), H.iptr(s.cols), H.dptr(s.vals), H.dptr(v), H.dptr(want))
    assert np.array_equal(_bits(_take(out, s.n)), _bits(want)), s.name
    gpu_lib.cgb200_invalidate(C.byref(A))


def test_mv_mult_ragged_rows_bit_exact(pkg, gpu_lib):
    # empty rows, single-entry rows and long rows mixed; sorted unique columns per row
    rng = np.random.default_rng(5)
    n = 5000
    rp, ci = [0], []
    for i in range(n):
        k = int(rng.choice([0, 1, 3, 7, 40, 900], p=[0.1, 0.2, 0.3, 0.3, 0.08, 0.02]))
        c = np.sort(rng.choice(n, size=min(k, n), replace=False))
        ci.extend(c.tolist())
        rp.append(len(ci))
    va = rng.standard_normal(len(ci))
    s = H.System(rp, ci, va, np.zeros(n), "ragged")
    A, _ = _structs(pkg, s)
    v = H.rhs(n, 98)
    out = pkg.MvSparseP()
    assert gpu_lib.mv_mult(C.byref(A), C.byref(_vec(pkg, v)), C.byref(out)) == 0, pkg.last_error()
    want = np.empty(n)
    H.oracle().oracle_spmv(n, H.iptr(s.row_ptr), H.iptr(s.cols), H.dptr(s.vals), H.dptr(v), H.dptr(want))
    assert np.array_equal(_bits(_take(out, n)), _bits(want))
    gpu_lib.cgb200_invalidate(C.byref(A))


def test_mv_mult_long_ragged_rows_take_the_long_row_kernel_bit_exact(pkg, gpu_lib):
    # mean row length >= 64 on one GPU: a long-row kernel (k_spmv_rowblock by default: 16 rows per
    # warp, products through a shared-memory tile, added in ascending column order).  Empty rows,
    # lengths around the slab width (32) and rows of thousands of entries, all mixed
    rng = np.random.default_rng(15)
    n = 6000
    rp, ci = [0], []
    for i in range(n):
        k = int(rng.choice([0, 1, 7, 8, 9, 31, 32, 33, 64, 345, 2500], p=[0.05, 0.05, 0.05, 0.05, 0.05, 0.05, 0.1, 0.1, 0.1, 0.3, 0.1]))
        c = np.sort(rng.choice(n, size=min(k, n), replace=False))
        ci.extend(c.tolist())
        rp.append(len(ci))
    assert len(ci) >= 128 * n                                       # the selection threshold (kernels.cu)
    va = rng.standard_normal(len(ci))
    s = H.System(rp, ci, va, np.zeros(n), "long_ragged")
    A, _ = _structs(pkg, s)
    want = np.empty(n)
    for seed in (97, 98):
        v = H.rhs(n, seed)
        vv = _vec(pkg, v)
        out = pkg.MvSparseP()
        assert gpu_lib.mv_mult(C.byref(A), C.byref(vv), C.byref(out)) == 0, pkg.last_error()
        H.oracle().oracle_spmv(n, H.iptr(s.row_ptr), H.iptr(s.cols), H.dptr(s.vals), H.dptr(v), H.dptr(want))
        assert np.array_equal(_bits(_take(out, n)), _bits(want))
    gpu_lib.cgb200_invalidate(C.byref(A))


@pytest.mark.parametrize("variant", ["classic", "sr"])
def test_long_row_system_solves_match_oracle(pkg, gpu_lib, variant):
    # a scaled-down course matrix (banded, 201 entries per row): fused p.s of the group-per-row kernel
    s = H.banded(4000, 100)
    d = _dev(pkg, s)
    if variant == "classic":
        xo, it, _, _ = H.oracle_cg(s, 12)
        r = d.solve(12)
    else:
        xo, it, _, _ = H.oracle_cg_sr(s, 12)
        r = d.solve(12, variant=pkg.SINGLE_REDUCTION)
    x = d.x()
    assert r.iterations == it
    assert np.linalg.norm(x - xo) <= TOL_X * np.linalg.norm(xo)
    v = H.rhs(s.n, 3)
    want = np.empty(s.n)
    H.oracle().oracle_spmv(s.n, H.iptr(s.row_ptr), H.iptr(s.cols), H.dptr(s.vals), H.dptr(v), H.dptr(want))
    assert np.array_equal(_bits(d.matvec(v)), _bits(want))
    d.free()


def _random_spd(rng, n, avg, symmetric=True):
    """Random sparse symmetric strictly diagonally dominant matrix (SPD), sorted unique columns per
    row, irreducible through a chain i -- i+1 (so the reference's row scan cannot bleed)."""
    nnz_off = max(1, int(avg * n / 2))
    i = rng.integers(0, n, nnz_off)
    j = rng.integers(0, n, nnz_off)
    keep = i != j
    i, j = i[keep], j[keep]
    chain = np.arange(n - 1)
    i = np.concatenate([i, chain])
    j = np.concatenate([j, chain + 1])
    if n >= 7000:  # two hub rows longer than one pipeline stage of the stream kernel (3584 entries)
        for hub in (3, n // 2):
            nb = rng.choice(n, size=4200, replace=False)
            nb = nb[nb != hub]
            i = np.concatenate([i, np.full(len(nb), hub)])
            j = np.concatenate([j, nb])
    lo, hi = np.minimum(i, j), np.maximum(i, j)     # one value per unordered pair, then mirror
    key = lo.astype(np.int64) * n + hi
    _, first = np.unique(key, return_index=True)
    lo, hi = lo[first], hi[first]
    v = rng.uniform(-1.0, 1.0, len(lo))
    rows = np.concatenate([lo, hi])
    cols = np.concatenate([hi, lo])
    vals = np.concatenate([v, v])
    absrow = np.zeros(n)
    np.add.at(absrow, rows, np.abs(vals))
    rows = np.concatenate([rows, np.arange(n)])
    cols = np.concatenate([cols, np.arange(n)])
    vals = np.concatenate([vals, absrow + 1.0])
    order = np.argsort(rows.astype(np.int64) * n + cols, kind="stable")
    rows, cols, vals = rows[order], cols[order], vals[order]
    rp = np.zeros(n + 1, dtype=np.int64)
    np.add.at(rp, rows + 1, 1)
    rp = np.cumsum(rp)
    return H.System(rp, cols, vals, H.rhs(n, int(rng.integers(1, 1 << 30))), f"random_spd_{n}_{avg}")


@pytest.mark.parametrize("seed", range(12))
def test_random_sparse_spd_systems_matvec_bits_and_fixed_count_solve(pkg, gpu_lib, seed):
    # randomised structure: sizes off every tile / vector-width boundary, row lengths from 1 to
    # hundreds (rows longer than a pipeline stage included at the larger densities)
    rng = np.random.default_rng(1000 + seed)
    n = int(rng.choice([1, 2, 5, 31, 511, 513, 1000, 2049, 4097, 7777, 20011]))
    avg = float(rng.choice([1.5, 4.0, 9.0, 30.0])) if n > 8 else 1.5
    s = _random_spd(rng, n, avg)
    A, _ = _structs(pkg, s)
    v = H.rhs(n, 55 + seed)
    out = pkg.MvSparseP()
    assert gpu_lib.mv_mult(C.byref(A), C.byref(_vec(pkg, v)), C.byref(out)) == 0, pkg.last_error()
    want = np.empty(n)
    H.oracle().oracle_spmv(n, H.iptr(s.row_ptr), H.iptr(s.cols), H.dptr(s.vals), H.dptr(v), H.dptr(want))
    full = s.nnz == n * n                      # tiny cases can come out dense: tree-summed path
    if full:
        assert np.allclose(_take(out, n), want, rtol=1e-13, atol=0)
    else:
        assert np.array_equal(_bits(_take(out, n)), _bits(want)), s.name
    gpu_lib.cgb200_invalidate(C.byref(A))
    M = int(rng.integers(0, 9))
    xo, it, _, st = H.oracle_cg(s, M, guard=1)
    for variant, orc in ((0, (xo, it)), (1, None)):
        x, r = gpu_solve(pkg, gpu_lib, s, M, 0.0, variant)
        if variant == 1:
            xs, its, _, _ = H.oracle_cg_sr(s, M)
            orc = (xs, its)
        if r.status in (pkg.EXACT, pkg.BREAKDOWN):
            continue                            # converged to round-off before M (tiny systems)
        assert r.iterations == orc[1], (s.name, variant)
        # the single-reduction recurrence amplifies rounding differences on badly scaled systems (the
        # hub rows): its two CPU forms already differ by `drift`, so allow the device that much
        drift = np.linalg.norm(orc[0] - xo) / np.linalg.norm(xo) if variant == 1 else 0.0
        err = np.linalg.norm(x - orc[0]) / np.linalg.norm(orc[0])
        assert err <= max(TOL_X, 20 * drift), (s.name, variant, err, drift)


@pytest.mark.parametrize("n", [1, 2, 3, 7, 64, 513, 1024, 2050, 2052, 4104])
def test_mv_mult_dense_path(pkg, gpu_lib, n):
    # full rows take the values-only GEMV; tree summation => tolerance, not bits.  n % 4 == 0 takes
    # the balanced unit-range kernel (2052 and 4104: row groups shared between CTAs, fixed up by the
    # last CTA), the rest the whole-row-group kernel
    s = H.dense(n)
    A, _ = _structs(pkg, s)
    v = H.rhs(n, 97)
    out = pkg.MvSparseP()
    assert gpu_lib.mv_mult(C.byref(A), C.byref(_vec(pkg, v)), C.byref(out)) == 0, pkg.last_error()
    want = np.empty(n)
    H.oracle().oracle_spmv(n, H.iptr(s.row_ptr), H.iptr(s.cols), H.dptr(s.vals), H.dptr(v), H.dptr(want))
    got = _take(out, n)
    scale = np.abs(s.vals.reshape(n, n)) @ np.abs(v)
    assert np.all(np.abs(got - want) <= 4 * n * EPS * scale + 1e-300)
    gpu_lib.cgb200_invalidate(C.byref(A))


# ---------------------------------------------------------------------------------------------
# the loop, fixed count (the reference's semantics): compare at equal K
# ---------------------------------------------------------------------------------------------

LOOP_CASES = [
    (lambda: H.lap(32, 32), 30),       # config 1 stand-in: ./cg file 30
    (lambda: H.lap(64, 64), 100),
    (lambda: H.lap(12, 12, 12), 25),
    (lambda: H.lap(40, 40, 40), 60),
    (lambda: H.banded(3000, 5), 12),
    (lambda: H.dense(512), 3),
    (lambda: H.dense(1000), 4),
    (lambda: H.dense(2052), 3),
    (lambda: H.lap(7, 5), 0),
    (lambda: H.lap(301, 299), 200),
]


@pytest.mark.parametrize("mk,M", LOOP_CASES)
def test_conj_grad_fixed_count_matches_oracle(pkg, gpu_lib, mk, M):
    s = mk()
    x = gpu_cg(pkg, gpu_lib, s, M)
    xo, it, _, _ = H.oracle_cg(s, M)
    assert it == M + 1
    rel = np.linalg.norm(x - xo) / np.linalg.norm(xo)
    assert rel <= TOL_X, (s.name, rel)


def test_conj_grad_is_deterministic(pkg, gpu_lib):
    s = H.lap(128, 128)
    x1 = gpu_cg(pkg, gpu_lib, s, 150)
    x2 = gpu_cg(pkg, gpu_lib, s, 150)
    assert np.array_equal(_bits(x1), _bits(x2))


def test_iteration_count_semantics(pkg, gpu_lib):
    # max_iter = M performs M+1 updates (cg.c:110-135)
    s = H.lap(20, 20)
    for M in (0, 1, 5, 33):
        _, r = gpu_solve(pkg, gpu_lib, s, M, 0.0)
        assert r.iterations == M + 1 and r.status == pkg.MAXITER


# ---------------------------------------------------------------------------------------------
# tolerance stop (extension): iteration count +-2, true residual, solution difference
# ---------------------------------------------------------------------------------------------

TOL_CASES = [
    lambda: H.lap(64, 64),
    lambda: H.lap(200, 150),
    lambda: H.lap(24, 24, 24),
    lambda: H.banded(4000, 8),
    lambda: H.dense(768),
]


@pytest.mark.parametrize("mk", TOL_CASES)
def test_tolerance_mode_matches_oracle(pkg, gpu_lib, mk):
    s = mk()
    x, r = gpu_solve(pkg, gpu_lib, s, 100000, TOL_RES)
    xo, it, rr, st = H.oracle_cg(s, 100000, tol=TOL_RES, guard=1)
    assert st == 1 and r.status == pkg.CONVERGED
    assert abs(r.iterations - it) <= ITER_SLACK, (r.iterations, it)
    assert r.relres <= TOL_RES
    assert H.true_relres(s, x) <= TOL_RES * 1.05
    assert np.linalg.norm(x - xo) / np.linalg.norm(xo) <= TOL_X


# ---------------------------------------------------------------------------------------------
# single-reduction variant (extension; the reference has one recurrence): against its own CPU
# restatement at equal K, and against the classic oracle (same x in exact arithmetic)
# ---------------------------------------------------------------------------------------------


@pytest.mark.parametrize("mk,M", LOOP_CASES)
def test_single_reduction_fixed_count(pkg, gpu_lib, mk, M):
    s = mk()
    x, r = gpu_solve(pkg, gpu_lib, s, M, 0.0, variant=pkg.SINGLE_REDUCTION)
    xs, it, _, _ = H.oracle_cg_sr(s, M)
    assert r.iterations == it == M + 1 and r.status == pkg.MAXITER
    assert np.linalg.norm(x - xs) / np.linalg.norm(xs) <= TOL_X, s.name
    xo, _, _, _ = H.oracle_cg(s, M)
    assert np.linalg.norm(x - xo) / np.linalg.norm(xo) <= TOL_X, s.name


@pytest.mark.parametrize("mk", TOL_CASES)
def test_single_reduction_tolerance(pkg, gpu_lib, mk):
    s = mk()
    x, r = gpu_solve(pkg, gpu_lib, s, 100000, TOL_RES, variant=pkg.SINGLE_REDUCTION)
    xo, it, rr, st = H.oracle_cg(s, 100000, tol=TOL_RES, guard=1)
    assert st == 1 and r.status == pkg.CONVERGED
    assert abs(r.iterations - it) <= ITER_SLACK, (r.iterations, it)
    assert H.true_relres(s, x) <= TOL_RES * 1.05
    assert np.linalg.norm(x - xo) / np.linalg.norm(xo) <= TOL_X


def test_single_reduction_known_answers_and_breakdown(pkg, gpu_lib):
    s = H.tridiag4()
    x, r = gpu_solve(pkg, gpu_lib, s, 1, 0.0, variant=pkg.SINGLE_REDUCTION)
    assert np.allclose(x, [2.0, 3.0, 3.0, 2.0], rtol=0, atol=1e-14) and r.iterations == 2
    # past exact convergence: stops (EXACT or BREAKDOWN on a residual of rounding noise), finite x
    x, r = gpu_solve(pkg, gpu_lib, s, 10, 0.0, variant=pkg.SINGLE_REDUCTION)
    assert np.all(np.isfinite(x)) and np.allclose(x, [2.0, 3.0, 3.0, 2.0], rtol=0, atol=1e-12)
    assert r.status in (pkg.EXACT, pkg.BREAKDOWN, pkg.MAXITER)
    # deterministic run to run
    t = H.lap(128, 128)
    x1, _ = gpu_solve(pkg, gpu_lib, t, 150, 0.0, variant=pkg.SINGLE_REDUCTION)
    x2, _ = gpu_solve(pkg, gpu_lib, t, 150, 0.0, variant=pkg.SINGLE_REDUCTION)
    assert np.array_equal(_bits(x1), _bits(x2))


def test_queued_fixed_count_solves_match_synchronous_ones(pkg, gpu_lib):
    # cgb200_system_solve_async: several right-hand sides queued back to back, collected later
    for variant in (pkg.CLASSIC, pkg.SINGLE_REDUCTION):
        d = pkg.DeviceSystem.lap(48, 40, 6)
        want = []
        for seed in (101, 102, 103):
            d.gen_rhs(seed)
            r = d.solve(25, variant=variant)
            want.append((r.iterations, r.rr, d.x()))
        tickets = []
        xs = []
        for seed in (101, 102, 103):
            d.gen_rhs(seed)
            tickets.append(d.solve_async(25, variant=variant))
        for i, t in enumerate(tickets):
            r = d.solve_wait(t)
            assert r.iterations == want[i][0] == 26 and r.status == pkg.MAXITER
            assert r.rr == want[i][1]          # same kernels, same order: same bits
        assert np.array_equal(_bits(d.x()), _bits(want[2][2]))
        # a tolerance cannot be queued, and a ticket is collected once
        o = pkg.Opts(10, 1e-8, 0, 0, 0)
        tk = C.c_longlong(-1)
        assert gpu_lib.cgb200_system_solve_async(d.h, C.byref(o), C.byref(tk)) == -1
        assert gpu_lib.cgb200_solve_wait(tickets[0], C.byref(pkg.Result())) == -1
        d.free()


def _dev(pkg, s):
    return pkg.DeviceSystem.from_csr(s.n, 0, s.row_ptr, s.cols, s.vals, s.b)


@pytest.mark.parametrize("make", [lambda: H.lap(64, 48), lambda: H.lap(37, 29), lambda: H.lap(31, 21, 17),
                                  lambda: H.banded(1201, 60), lambda: H.banded(3002, 40), lambda: H.lap(1, 1),
                                  lambda: H.lap(3, 1), lambda: H.lap(500, 600)],
                         ids=["lap64x48", "lap37x29_n%4=1", "lap3d_n%4=3", "banded1201", "banded3002_n%4=2", "n=1", "n=3",
                              "lap500x600"])
def test_fused_vector_update_has_the_bits_of_the_three_kernel_loop(pkg, gpu_lib, make):
    # small single-GPU systems: K2 and K3 as one co-operative launch (k_update_xrp).  Same element-to-
    # thread mapping and summation order as k_update_xr, so x, r.r and the iteration count are
    # bit-identical to the three-kernel loop; and both agree with the oracle
    s = make()
    xo = H.oracle_cg(s, 12, guard=1)[0]            # (guard: stop where the reference would divide by r.r = 0)
    got = {}
    try:
        for fused in (1, 0):
            gpu_lib.cgb200_set_fused_vector_update(fused)
            d = _dev(pkg, s)
            r = d.solve(12, check_every=5)
            x12 = d.x()
            rt = d.solve(5000, tol=1e-11)
            got[fused] = (x12, r.rr, r.iterations, r.kernel_launches, d.x(), rt.iterations, rt.status, rt.rr)
            d.free()
    finally:
        gpu_lib.cgb200_set_fused_vector_update(1)
    f, u = got[1], got[0]
    assert f[2] == u[2] and f[5] == u[5] and f[6] == u[6]
    assert np.array_equal(f[0].view(np.uint64), u[0].view(np.uint64))
    assert np.array_equal(f[4].view(np.uint64), u[4].view(np.uint64))
    assert f[1] == u[1] and f[7] == u[7]
    if u[2] == 13:                                   # (tiny systems stop early: exact solution reached)
        assert f[3] == 2 * 13 + 1 and u[3] == 3 * 13 + 1
    assert np.linalg.norm(f[0] - xo) <= 1e-8 * max(np.linalg.norm(xo), 1e-300)


@pytest.mark.parametrize("n", [256, 1000, 2052, 4096])
def test_dense_one_kernel_iteration_matches_three_kernel_form(pkg, gpu_lib, n):
    # dense path: loop entry and every iteration in ONE cooperative launch per solve
    # (k_cg_dense_solve) against the K1/K2/K3 form of the same recurrence, and against the oracle
    s = H.dense(n)
    xo, it_o, _, _ = H.oracle_cg(s, 4)
    got = {}
    try:
        for on in (1, 0):
            gpu_lib.cgb200_set_one_kernel_iteration(on)
            d = _dev(pkg, s)
            assert d.one_kernel_iteration() == bool(on)
            l0 = gpu_lib.cgb200_launch_count()
            r = d.solve(4, check_every=5)
            launches = gpu_lib.cgb200_launch_count() - l0
            assert r.iterations == it_o == 5 and r.status == pkg.MAXITER
            assert launches == (1 if on else 1 + 10), launches      # one launch per solve | k_init + 5 x (K1, K2+K3 fused: small system)
            rt = d.solve(500, tol=1e-10)
            got[on] = (d.x(), r.rr, rt.iterations, rt.status, d.true_relres())
            d.gen_rhs(3)
            r1 = d.solve(4)
            x1 = d.x()
            d.gen_rhs(3)
            r2 = d.solve(4)
            assert r1.rr == r2.rr and np.array_equal(_bits(x1), _bits(d.x()))   # run-to-run bits
            d.free()
    finally:
        gpu_lib.cgb200_set_one_kernel_iteration(1)
    for on in (0, 1):
        assert got[on][3] == pkg.CONVERGED and got[on][4] <= TOL_RES * 1.05
    assert abs(got[1][2] - got[0][2]) <= 1
    assert np.linalg.norm(got[1][0] - got[0][0]) <= 1e-12 * np.linalg.norm(got[0][0])
    assert abs(got[1][1] - got[0][1]) <= 1e-10 * abs(got[0][1])
    d = _dev(pkg, s)
    d.solve(4)
    assert np.linalg.norm(d.x() - xo) <= TOL_X * np.linalg.norm(xo)
    d.free()


def test_dense_one_kernel_iteration_breakdown_and_exact_stop(pkg, gpu_lib):
    # the stop flag raised behind a grid barrier must stop every CTA of the launch: identity
    # matrix converges exactly after one update (status EXACT, x = b), zero rhs stops at entry
    n = 64
    s = H.dense(n)
    s.vals[:] = 0.0
    s.vals[::n + 1] = 1.0
    d = _dev(pkg, s)
    assert d.one_kernel_iteration()
    r = d.solve(10)
    assert r.status == pkg.EXACT and r.iterations == 1
    assert np.array_equal(d.x(), s.b)
    d.set_rhs(np.zeros(n))
    r = d.solve(10)
    assert r.status == pkg.EXACT and r.iterations == 0 and not np.any(d.x())
    d.free()


# ---------------------------------------------------------------------------------------------
# resident mode of the operator API: the reference's loop, statement by statement
# ---------------------------------------------------------------------------------------------


@pytest.mark.parametrize("mk,M", [(lambda: H.lap(32, 32), 30), (lambda: H.lap(9, 7, 5), 10), (lambda: H.dense(257), 4)])
def test_reference_loop_through_the_operators_in_resident_mode(pkg, gpu_lib, mk, M):
    # cg.c:104-135 written against mv_ops.h, as a caller of the library would: in resident mode
    # nothing but the dot products' return values crosses the bus until cgb200_to_host
    L = gpu_lib
    s = mk()
    A, b = _structs(pkg, s)
    xo, it, _, _ = H.oracle_cg(s, M)
    assert L.cgb200_set_mode(pkg.RESIDENT) == 0
    try:
        x_new = L.new_mv_struct_with_size(s.n)
        r = L.mv_deep_copy(C.byref(b))
        p = L.mv_deep_copy(r)
        sv, sres = pkg.MvSparseP(), pkg.MvSparseP()
        k = 0
        while True:
            assert L.mv_mult(C.byref(A), p, C.byref(sv)) == 0, pkg.last_error()
            alpha = L.dot_product(r, r) / L.dot_product(p, sv)
            assert L.sv_mult(alpha, p, C.byref(sres)) == 0
            x_old = L.mv_deep_copy(x_new)                         # device to device
            assert L.vec_add(x_old, sres, C.byref(x_new)) == 0
            r_prev = L.mv_deep_copy(r)
            assert L.sv_mult(alpha, sv, C.byref(sres)) == 0
            assert L.vec_sub(r, sres, C.byref(r)) == 0             # in place (cg.c:123)
            if k == M:
                break
            beta = L.dot_product(r, r) / L.dot_product(r_prev, r_prev)
            assert L.sv_mult(beta, p, C.byref(sres)) == 0
            assert L.vec_add(r, sres, C.byref(p)) == 0
            k += 1
            for o in (x_old, r_prev):
                L.cgb200_free_all(o)
        stale = _take(x_new, s.n)
        assert L.cgb200_to_host(x_new) == 0
        x = _take(x_new, s.n)
    finally:
        assert L.cgb200_set_mode(pkg.EAGER) == 0
        L.cgb200_invalidate(C.byref(A))
    assert not np.any(stale)                  # the host array was never written before to_host
    assert np.linalg.norm(x - xo) <= TOL_X * np.linalg.norm(xo)


@pytest.mark.gpu
def test_resident_mode_stored_self_dot_follows_every_write(pkg, gpu_lib):
    # resident mode answers a repeated dot_product(v, v) from the value stored with v (the
    # reference's loop asks for r.r twice per residual): every write must drop it, a deep copy
    # carries it, and the answer is bit-identical to the kernel's in eager mode
    L = gpu_lib
    n = 5000
    a = H.rhs(n, 11)
    c = H.rhs(n, 12)
    va, vc = _vec(pkg, a), _vec(pkg, c)
    eager = L.dot_product(C.byref(va), C.byref(va))
    s = a + c
    vs = _vec(pkg, s)
    eager_sum = L.dot_product(C.byref(vs), C.byref(vs))
    assert L.cgb200_set_mode(pkg.RESIDENT) == 0
    try:
        r = L.mv_deep_copy(C.byref(va))
        d1 = L.dot_product(r, r)
        d2 = L.dot_product(r, r)                                   # stored value
        cp = L.mv_deep_copy(r)
        d3 = L.dot_product(cp, cp)                                 # carried by the copy
        assert L.vec_add(r, C.byref(vc), C.byref(r)) == 0           # in-place write
        d4 = L.dot_product(r, r)
        d5 = L.dot_product(cp, cp)                                 # the copy did not change
        assert L.sv_mult(2.0, r, C.byref(cp)) == 0                  # cp overwritten
        d6 = L.dot_product(cp, cp)
        for o in (r, cp):
            L.cgb200_free_all(o)
    finally:
        assert L.cgb200_set_mode(pkg.EAGER) == 0
        for v in (va, vc, vs):
            L.cgb200_invalidate(C.byref(v))
    assert _bits(np.array([d1, d2, d3, d5])).tolist() == [_bits(np.array([eager]))[0]] * 4
    assert _bits(np.array([d4]))[0] == _bits(np.array([eager_sum]))[0]
    assert d6 == 4.0 * eager_sum                                    # scaling by 2 is exact


@pytest.mark.gpu
@pytest.mark.parametrize("n", [1, 7, 4096, 50001])
def test_resident_mode_deferred_scale_is_invisible(pkg, gpu_lib, n):
    # resident mode defers sv_mult and lets the vec_add / vec_sub that reads the result form
    # a +/- fl(s*b) in one pass; every other reader, and every write to the source, must see exactly
    # what the eager operators produce (numpy: the same unfused roundings)
    L = gpu_lib
    a, c = H.rhs(n, 21), H.rhs(n, 22)
    va, vc = _vec(pkg, a), _vec(pkg, c)
    s1, s2 = 0.3721, -1.9

    def host(ptr):
        assert L.cgb200_to_host(ptr) == 0
        return _take(ptr, n)

    got = {}
    assert L.cgb200_set_mode(pkg.RESIDENT) == 0
    try:
        ra, rc = L.mv_deep_copy(C.byref(va)), L.mv_deep_copy(C.byref(vc))
        t, u, w = pkg.MvSparseP(), pkg.MvSparseP(), pkg.MvSparseP()
        assert L.sv_mult(s1, ra, C.byref(t)) == 0                   # deferred
        assert L.vec_add(rc, t, C.byref(u)) == 0                    # consumed in one pass
        got["add"] = host(u)
        got["t_after_use"] = host(t)                                # ... and still readable
        assert L.sv_mult(s1, ra, C.byref(t)) == 0
        assert L.vec_sub(rc, t, C.byref(t)) == 0                    # output is the deferred vector
        got["sub_into_t"] = host(t)
        assert L.sv_mult(s1, ra, C.byref(t)) == 0
        assert L.vec_add(ra, rc, C.byref(ra)) == 0                  # the source changes first
        got["t_before_src_write"] = host(t)
        got["ra_sum"] = host(ra)
        assert L.sv_mult(s2, ra, C.byref(t)) == 0
        assert L.vec_add(rc, t, C.byref(ra)) == 0                   # the output is the source
        got["out_is_src"] = host(ra)
        got["t_after_out_is_src"] = host(t)                         # stored by the same pass
        assert L.sv_mult(s1, rc, C.byref(t)) == 0
        assert L.sv_mult(s2, rc, C.byref(t)) == 0                   # overwritten unread
        assert L.sv_mult(s1, t, C.byref(w)) == 0                    # deferred on a deferred vector
        assert L.sv_mult(s1, rc, C.byref(t)) == 0                   # its source is rewritten first
        got["chain"] = host(w)
        d = L.dot_product(t, t)                                     # another reader
        cp = L.mv_deep_copy(t)
        got["copy"] = host(cp)
        assert L.sv_mult(s2, rc, C.byref(t)) == 0
        L.cgb200_free_all(rc)                                       # the source goes away
        got["src_freed"] = host(t)
        assert L.sv_mult(s1, ra, C.byref(t)) == 0
        assert L.vec_sub(ra, t, C.byref(ra)) == 0                   # in place on the first operand
        got["in_place"] = host(ra)
        assert L.sv_mult(s1, ra, C.byref(t)) == 0                   # left deferred at the mode switch
    finally:
        assert L.cgb200_set_mode(pkg.EAGER) == 0
    got["left_deferred"] = _take(t, n)
    for o in (ra, t, u, w, cp):
        L.cgb200_free_all(o)
    for v in (va, vc):
        L.cgb200_invalidate(C.byref(v))
    sa = s1 * a
    want = {"add": c + sa, "t_after_use": sa, "sub_into_t": c - sa, "t_before_src_write": sa, "ra_sum": a + c}
    ra1 = a + c
    want["out_is_src"] = c + s2 * ra1
    ra2 = want["out_is_src"]
    want["t_after_out_is_src"] = s2 * ra1
    want["chain"] = s1 * (s2 * c)
    want["copy"] = s1 * c
    want["src_freed"] = s2 * c
    want["in_place"] = ra2 - s1 * ra2
    want["left_deferred"] = s1 * want["in_place"]
    for k, v in want.items():
        assert np.array_equal(_bits(got[k]), _bits(v)), k
    ref = _vec(pkg, want["copy"])                                   # eager dot of the same values
    assert _bits(np.array([d]))[0] == _bits(np.array([L.dot_product(C.byref(ref), C.byref(ref))]))[0]


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["eager", "resident"])
@pytest.mark.parametrize("mk", [lambda: H.lap(40, 23), lambda: H.dense(130), lambda: H.lap(9, 8, 7)])
def test_mv_mult_writes_straight_into_the_output_vector(pkg, gpu_lib, mk, mode):
    # the operator's matvec gathers from the caller's device vector and stores into the output's
    # device buffer (no staging); an existing output is reused, and an output that IS the input
    # (not something the reference supports, mv_ops.c:178-181 zeroes it first) must not corrupt
    L = gpu_lib
    s = mk()
    A, _ = _structs(pkg, s)
    v = H.rhs(s.n, 5)
    want = np.empty(s.n)
    H.oracle().oracle_spmv(s.n, H.iptr(s.row_ptr), H.iptr(s.cols), H.dptr(s.vals), H.dptr(v), H.dptr(want))
    vhost = v.copy()                                              # must outlive the struct that points at it
    vv = _vec(pkg, vhost)
    assert L.cgb200_set_mode(pkg.RESIDENT if mode == "resident" else pkg.EAGER) == 0
    try:
        out = pkg.MvSparseP()
        assert L.mv_mult(C.byref(A), C.byref(vv), C.byref(out)) == 0, pkg.last_error()
        assert L.mv_mult(C.byref(A), C.byref(vv), C.byref(out)) == 0      # existing output reused
        assert L.cgb200_to_host(out) == 0
        got = _take(out, s.n)
        w = L.mv_deep_copy(C.byref(vv))
        assert L.mv_mult(C.byref(A), w, C.byref(w)) == 0                  # output aliases the input
        assert L.cgb200_to_host(w) == 0
        alias = _take(w, s.n)
        for o in (out, w):
            L.cgb200_free_all(o)
    finally:
        assert L.cgb200_set_mode(pkg.EAGER) == 0
        L.cgb200_invalidate(C.byref(A))
        L.cgb200_invalidate(C.byref(vv))
    if s.nnz == s.n * s.n:     # dense path: row sums are tree-ordered, not the reference's sequence
        scale = 4 * s.n * np.finfo(float).eps * (np.abs(s.vals).reshape(s.n, s.n) @ np.abs(v))
        assert np.all(np.abs(got - want) <= scale) and np.array_equal(_bits(alias), _bits(got))
    else:
        assert np.array_equal(_bits(got), _bits(want))
        assert np.array_equal(_bits(alias), _bits(want))


# ---------------------------------------------------------------------------------------------
# row-pattern dictionary format: same arithmetic, one byte per row
# ---------------------------------------------------------------------------------------------

PATTERN_SYSTEMS = [
    lambda: H.tridiag4(),
    lambda: H.lap(32, 32),
    lambda: H.lap(100, 37),
    lambda: H.lap(12, 12, 12),
    lambda: H.lap(33, 17, 9),
    lambda: H.lap(1, 1, 700),                                     # 1D chain
    lambda: H.toeplitz(3000, [9.0, -1.0, -0.5, 0.25, -2.0]),
    lambda: H.toeplitz(1500, [12.0] + [-0.5 + 0.03 * k for k in range(1, 14)]),   # 27 entries per row: 4 passes of 8
    lambda: H.lap(256, 256),
]


@pytest.mark.parametrize("mk", PATTERN_SYSTEMS)
def test_pattern_format_matvec_bit_exact_and_solves(pkg, gpu_lib, mk):
    s = mk()
    d = _dev(pkg, s)
    v = H.rhs(s.n, 99)
    want = np.empty(s.n)
    H.oracle().oracle_spmv(s.n, H.iptr(s.row_ptr), H.iptr(s.cols), H.dptr(s.vals), H.dptr(v), H.dptr(want))
    csr_mv = d.matvec(v)
    M = 1 if s.n <= 4 else 12
    r0 = d.solve(M)
    x_csr = d.x()
    assert d.use_pattern_format(True), pkg.last_error()
    it_b, k1_b = d.bytes()
    assert it_b == s.n + 88 * s.n and k1_b == s.n + 16 * s.n        # one byte per row
    pat_mv = d.matvec(v)
    assert np.array_equal(_bits(pat_mv), _bits(want)) and np.array_equal(_bits(csr_mv), _bits(want)), s.name
    for variant in (pkg.CLASSIC, pkg.SINGLE_REDUCTION):
        r1 = d.solve(M, variant=variant)
        assert r1.iterations == r0.iterations and r1.status == r0.status
        x_pat = d.x()
        assert np.linalg.norm(x_pat - x_csr) <= 1e-11 * np.linalg.norm(x_csr), (s.name, variant)
    xo, it_o, _, _ = H.oracle_cg(s, M)
    d.solve(M)
    assert np.linalg.norm(d.x() - xo) <= TOL_X * np.linalg.norm(xo)
    # and back: the CSR kernel again, same bits as before the switch
    assert d.use_pattern_format(False) is False
    d.solve(M)
    assert np.array_equal(_bits(d.x()), _bits(x_csr))
    d.free()


def test_pattern_format_refuses_matrices_without_few_patterns(pkg, gpu_lib):
    s = H.banded(3000, 5)                # random values: every row is its own pattern
    d = _dev(pkg, s)
    assert d.use_pattern_format(True) is False
    assert "pattern format not applicable" in pkg.last_error()
    xo, _, _, _ = H.oracle_cg(s, 10)
    d.solve(10)                           # still on CSR, still right
    assert np.linalg.norm(d.x() - xo) <= TOL_X * np.linalg.norm(xo)
    d.free()
    d = pkg.DeviceSystem.dense(256)       # dense path has no index arrays to compress
    assert d.use_pattern_format(True) is False
    d.free()
    s = H.toeplitz(400, [60.0] + [-0.7] * 40)   # 81 entries per row: longer than a dictionary entry
    d = _dev(pkg, s)
    assert d.use_pattern_format(True) is False
    d.free()


def test_pattern_format_device_generated_laplacian_converges(pkg, gpu_lib):
    d = pkg.DeviceSystem.lap(96, 80, 48)
    r_csr = d.solve(5000, tol=1e-10)
    assert d.use_pattern_format(True)
    r_pat = d.solve(5000, tol=1e-10)
    assert r_pat.status == pkg.CONVERGED and abs(r_pat.iterations - r_csr.iterations) <= ITER_SLACK
    assert d.true_relres() <= TOL_RES * 1.05
    d.free()


# ---------------------------------------------------------------------------------------------
# device-resident systems and generators
# ---------------------------------------------------------------------------------------------


@pytest.mark.parametrize("dims", [(16, 16, 1), (37, 21, 1), (8, 8, 8), (19, 11, 7), (64, 64, 16)])
def test_device_generated_laplacian_matches_oracle_generator(pkg, gpu_lib, dims):
    nx, ny, nz = dims
    s = H.lap(nx, ny, nz)
    d = pkg.DeviceSystem.lap(nx, ny, nz)
    assert d.rows == s.n and d.nnz == s.nnz
    v = H.rhs(s.n, 55)
    want = np.empty(s.n)
    H.oracle().oracle_spmv(s.n, H.iptr(s.row_ptr), H.iptr(s.cols), H.dptr(s.vals), H.dptr(v), H.dptr(want))
    assert np.array_equal(_bits(d.matvec(v)), _bits(want))
    # same right-hand side generator => same solve as the host-struct path
    r = d.solve(20)
    xo, it, _, _ = H.oracle_cg(s, 20)
    assert r.iterations == it
    assert np.linalg.norm(d.x() - xo) / np.linalg.norm(xo) <= TOL_X
    d.free()


def test_device_generated_dense_matches_oracle_generator(pkg, gpu_lib):
    n = 640
    s = H.dense(n)
    d = pkg.DeviceSystem.dense(n)
    # A e_j is column j, reproduced exactly (one nonzero product per row)
    for j in (0, 1, 317, n - 1):
        e = np.zeros(n)
        e[j] = 1.0
        assert np.array_equal(d.matvec(e), s.vals.reshape(n, n)[:, j])
    r = d.solve(100, tol=TOL_RES)
    xo, it, _, _ = H.oracle_cg(s, 100, tol=TOL_RES, guard=1)
    assert abs(r.iterations - it) <= ITER_SLACK
    assert np.linalg.norm(d.x() - xo) / np.linalg.norm(xo) <= TOL_X
    assert d.true_relres() <= TOL_RES * 1.05
    d.free()


# ---------------------------------------------------------------------------------------------
# full-size configurations (BASELINE.json configs 2-4) through size-independent properties
# ---------------------------------------------------------------------------------------------


def test_full_size_lap2d_4096(pkg, gpu_lib):
    nx = ny = 4096
    d = pkg.DeviceSystem.lap(nx, ny, 1)
    assert d.rows == nx * ny and d.nnz == 5 * nx * ny - 2 * nx - 2 * ny
    # A*1 counts the missing neighbours: 0 inside, 1 on edges, 2 in corners
    s = d.matvec(np.ones(nx * ny)).reshape(ny, nx)
    want = np.zeros((ny, nx))
    want[0, :] += 1; want[-1, :] += 1; want[:, 0] += 1; want[:, -1] += 1
    assert np.array_equal(s, want)
    # linearity: A(u + 2v) = Au + 2Av exactly for small-integer vectors
    rng = np.random.default_rng(0)
    u = rng.integers(-8, 9, nx * ny).astype(np.float64)
    v = rng.integers(-8, 9, nx * ny).astype(np.float64)
    assert np.array_equal(d.matvec(u + 2 * v), d.matvec(u) + 2 * d.matvec(v))
    # the loop: 300 iterations reduce the residual and the recursive residual tracks the true one
    r = d.solve(299)
    assert r.iterations == 300 and r.status == pkg.MAXITER
    tr = d.true_relres()
    assert tr < 0.2 and abs(tr - r.relres) <= 1e-6 * max(tr, r.relres)
    d.free()


def test_full_size_lap3d_512_slab_consistency(pkg, gpu_lib):
    # one 512x512x64 slab of the 512^3 operator (what one of 8 ranks owns) as a standalone system:
    # stencil sums and row counts
    d = pkg.DeviceSystem.lap(512, 512, 64)
    n = 512 * 512 * 64
    assert d.rows == n and d.nnz == 7 * n - 2 * (512 * 64 * 2) - 2 * 512 * 512
    s = d.matvec(np.ones(n)).reshape(64, 512, 512)
    want = np.zeros((64, 512, 512))
    want[0] += 1; want[-1] += 1; want[:, 0, :] += 1; want[:, -1, :] += 1; want[:, :, 0] += 1; want[:, :, -1] += 1
    assert np.array_equal(s, want)
    d.free()


def test_full_size_dense_32768(pkg, gpu_lib):
    n = 32768
    d = pkg.DeviceSystem.dense(n)
    O = H.oracle()
    # symmetry + generator parity through A e_j (exact), a few columns
    for j in (0, 12345, n - 1):
        e = np.zeros(n)
        e[j] = 1.0
        col = d.matvec(e)
        for i in (0, 1, 777, j, n - 1):
            want = 0.5 * n + 1.0 if i == j else O.oracle_uniform(1, min(i, j) * n + max(i, j))
            assert col[i] == want
    # strictly diagonally dominant => converges in a handful of iterations to 1e-10
    r = d.solve(100, tol=TOL_RES)
    assert r.status == pkg.CONVERGED and r.iterations <= 12
    assert d.true_relres() <= TOL_RES * 1.05
    d.free()


# ---------------------------------------------------------------------------------------------
# full-size configurations against the ORACLE itself (north_star: "results must match the
# reference's CPU solve on the same inputs"): x after a fixed small K at BASELINE sizes (SURVEY
# H9), the matvec bit for bit on all rows, and the device generator against the oracle's.
# ---------------------------------------------------------------------------------------------


def _host_ram_gb():
    try:
        import psutil

        return psutil.virtual_memory().available / 1e9
    except Exception:  # noqa: BLE001
        return 0.0


def test_full_size_lap2d_4096_x_matches_the_oracle_after_20_iterations(pkg, gpu_lib):
    L = gpu_lib
    nx = ny = 4096
    n = nx * ny
    s = H.lap(nx, ny)                                  # the oracle's own generator, 83.9 M entries
    d = pkg.DeviceSystem.lap(nx, ny, 1)
    # the device-generated system IS the oracle's, bit for bit
    vals = np.empty(d.nnz)
    cols = np.empty(d.nnz, dtype=np.int32)
    rp = np.empty(n + 1, dtype=np.int32)
    assert L.cgb200_system_download(d.h, H.dptr(vals), H.iptr(cols), H.iptr(rp)) == 0, pkg.last_error()
    assert np.array_equal(rp, s.row_ptr) and np.array_equal(cols, s.cols) and np.array_equal(_bits(vals), _bits(s.vals))
    del vals, cols, rp
    # the matvec on all 16.7 M rows, bit-exact
    v = H.rhs(n, 77)
    want = np.empty(n)
    H.oracle().oracle_spmv(n, H.iptr(s.row_ptr), H.iptr(s.cols), H.dptr(s.vals), H.dptr(v), H.dptr(want))
    assert np.array_equal(_bits(d.matvec(v)), _bits(want))
    # the loop: K = 20 iterations, device-resident system and the reference's own entry point
    K = 20
    xo, it, rr_o, _ = H.oracle_cg(s, K - 1)
    assert it == K
    r = d.solve(K - 1)
    assert r.iterations == K and r.status == pkg.MAXITER
    x = d.x()
    assert np.linalg.norm(x - xo) / np.linalg.norm(xo) <= TOL_X
    assert abs(r.rr - rr_o) <= 1e-10 * rr_o
    d.free()
    x2 = gpu_cg(pkg, L, s, K - 1)                      # conj_grad(19, A, b, &x) on the oracle's host arrays
    assert np.linalg.norm(x2 - xo) / np.linalg.norm(xo) <= TOL_X
    assert np.array_equal(_bits(x2), _bits(x))         # same kernels, same bits, whichever way A arrived


def test_full_size_lap3d_512_x_matches_the_oracle_after_5_iterations(pkg, gpu_lib):
    if _host_ram_gb() < 48:
        pytest.skip("needs ~25 GB of host memory for the 512^3 CSR system and the oracle's vectors")
    nx = 512
    n = nx ** 3
    s = H.lap(nx, nx, nx)                              # 938 M entries, 11.8 GB
    d = pkg.DeviceSystem.lap(nx, nx, nx)
    assert d.rows == n and d.nnz == s.nnz
    v = H.rhs(n, 78)
    want = np.empty(n)
    H.oracle().oracle_spmv(n, H.iptr(s.row_ptr), H.iptr(s.cols), H.dptr(s.vals), H.dptr(v), H.dptr(want))
    assert np.array_equal(_bits(d.matvec(v)), _bits(want))   # generator + kernel on all 134 M rows
    del v, want
    K = 5
    xo, it, rr_o, _ = H.oracle_cg(s, K - 1)
    assert it == K
    r = d.solve(K - 1)
    assert r.iterations == K and r.status == pkg.MAXITER
    x = d.x()
    assert np.linalg.norm(x - xo) / np.linalg.norm(xo) <= TOL_X
    assert abs(r.rr - rr_o) <= 1e-10 * rr_o
    if d.use_pattern_format(True):                     # the row-pattern format on the same system
        d.solve(K - 1)                                 # same row sums bit for bit; the dot products are
        xp = d.x()                                     # summed over another grid shape
        assert np.linalg.norm(xp - xo) / np.linalg.norm(xo) <= TOL_X
        assert np.linalg.norm(xp - x) / np.linalg.norm(x) <= 1e-12
    d.free()


def test_full_size_dense_32768_block_matches_the_oracle(pkg, gpu_lib):
    # the dense generator and GEMV against the oracle on a row block of the n = 32768 system that the
    # oracle can hold (rows 20000..20511): A[rows, :] p for a random p, within the dense path's bar
    n, r0, nr = 32768, 20000, 512
    O = H.oracle()
    d = pkg.DeviceSystem.dense(n)
    p = H.rhs(n, 79)
    got = d.matvec(p)[r0:r0 + nr]
    blk = np.empty((nr, n))
    for i in range(nr):
        gi = r0 + i
        for j in (0, gi - 1, gi, gi + 1, n - 1):
            blk[i, j] = 0.5 * n + 1.0 if j == gi else O.oracle_uniform(1, min(gi, j) * n + max(gi, j))
    # full rows through a vectorised restatement of the same hash (splitmix64, as oracle_uniform)
    def u(ctr):
        with np.errstate(over="ignore"):
            def sm64(z):
                z = z + np.uint64(0x9E3779B97F4A7C15)
                z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
                z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
                return z ^ (z >> np.uint64(31))
            h = sm64(sm64(np.uint64(1)) ^ (ctr * np.uint64(0xD1342543DE82EF95) + np.uint64(1)))
        return (h >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0) - 0.5
    j = np.arange(n, dtype=np.uint64)
    for i in range(nr):
        gi = np.uint64(r0 + i)
        lo, hi = np.minimum(gi, j), np.maximum(gi, j)
        row = u(lo * np.uint64(n) + hi)
        row[r0 + i] = 0.5 * n + 1.0
        for jj in (0, r0 + i - 1, r0 + i, r0 + i + 1, n - 1):   # the vectorised hash IS oracle_uniform
            assert row[jj] == blk[i, jj]
        blk[i] = row
    want = blk @ p
    scale = np.abs(blk) @ np.abs(p)
    assert np.all(np.abs(got - want) <= 4 * n * EPS * scale)
    d.free()

