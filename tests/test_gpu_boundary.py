"""GPU: behaviour at the drop-in boundary that an UNMODIFIED caller of mv_ops.h relies on --
in-place edits of a matrix between calls (the reference reads the caller's arrays on every call,
mv_ops.c:187-197), negative-definite systems (the reference divides by p.Ap whatever its sign,
cg.c:113), malformed CSR input, exact launch counts of fixed-count solves, and calls from several
threads (the reference's operators are stateless)."""
import ctypes as C
import threading

import numpy as np
import pytest

import helpers as H

pytestmark = pytest.mark.gpu

TOL_X = 1e-8


def _structs(pkg, sysm):
    A = pkg.MvSparse(sysm.n, sysm.nnz, H.dptr(sysm.vals), H.iptr(sysm.cols), H.iptr(sysm.row_ptr))
    b = pkg.MvSparse(sysm.n, sysm.n, H.dptr(sysm.b), None, None)
    return A, b


def _take(ptr, n):
    return np.ctypeslib.as_array(ptr.contents.values, shape=(n,)).copy()


def _bits(a):
    return np.ascontiguousarray(a).view(np.uint64)


def _spmv(sysm, v):
    out = np.empty(sysm.n)
    H.oracle().oracle_spmv(sysm.n, H.iptr(sysm.row_ptr), H.iptr(sysm.cols), H.dptr(sysm.vals), H.dptr(v), H.dptr(out))
    return out


def _off_grid_index(nnz):
    """An entry the 64-sample fingerprint (abi.cu: fingerprint) never looks at."""
    step = nnz // 64 if nnz > 64 else 1
    k = step // 2 + 3 * step
    assert k % step != 0 and k != nnz - 1
    return k


@pytest.fixture
def strict(pkg, gpu_lib):
    assert gpu_lib.cgb200_set_matrix_cache(pkg.CACHE_STRICT) == 0
    yield gpu_lib
    gpu_lib.cgb200_set_matrix_cache(pkg.CACHE_TRACKED)


def test_in_place_edit_between_two_mv_mult_calls_is_seen(pkg, strict):
    # default policy, no cgb200_invalidate: exactly what an unmodified caller does
    L = strict
    s = H.lap(40, 30)
    A, _ = _structs(pkg, s)
    v = H.rhs(s.n, 31)
    vv = pkg.MvSparse(s.n, s.n, H.dptr(v), None, None)
    out = pkg.MvSparseP()
    assert L.mv_mult(C.byref(A), C.byref(vv), C.byref(out)) == 0, pkg.last_error()
    assert np.array_equal(_bits(_take(out, s.n)), _bits(_spmv(s, v)))
    k = _off_grid_index(s.nnz)
    s.vals[k] = 123.25
    assert L.mv_mult(C.byref(A), C.byref(vv), C.byref(out)) == 0, pkg.last_error()
    assert np.array_equal(_bits(_take(out, s.n)), _bits(_spmv(s, v))), "stale device matrix"
    # a column index edit (same sparsity count) is seen too
    row = int(np.searchsorted(s.row_ptr, k, side="right") - 1)
    if s.row_ptr[row + 1] - 1 == k and s.cols[k] + 1 < s.n:   # last entry of its row: move it right
        s.cols[k] += 1
    else:
        k2 = s.row_ptr[row + 1] - 1
        if s.cols[k2] + 1 < s.n:
            s.cols[k2] += 1
    assert L.mv_mult(C.byref(A), C.byref(vv), C.byref(out)) == 0, pkg.last_error()
    assert np.array_equal(_bits(_take(out, s.n)), _bits(_spmv(s, v)))
    L.cgb200_invalidate(C.byref(A))
    L.cgb200_free_all(out)


def test_in_place_edit_between_two_conj_grad_calls_is_seen(pkg, strict):
    L = strict
    s = H.lap(24, 24)
    A, b = _structs(pkg, s)
    xp = pkg.MvSparseP()
    assert L.conj_grad(20, C.byref(A), C.byref(b), C.byref(xp)) == 0
    x1 = _take(xp, s.n)
    assert np.linalg.norm(x1 - H.oracle_cg(s, 20)[0]) <= TOL_X * np.linalg.norm(x1)
    s.vals[s.row_ptr[1:] - 1] *= 1.0   # no-op edit: still cached, same answer
    for i in range(s.n):               # scale the diagonal in place: a different (still SPD) matrix
        for k in range(s.row_ptr[i], s.row_ptr[i + 1]):
            if s.cols[k] == i:
                s.vals[k] = 4.5
    assert L.conj_grad(20, C.byref(A), C.byref(b), C.byref(xp)) == 0
    x2 = _take(xp, s.n)
    xo = H.oracle_cg(s, 20)[0]
    assert np.linalg.norm(x2 - xo) <= TOL_X * np.linalg.norm(xo)
    assert np.linalg.norm(x2 - x1) > 1e-3 * np.linalg.norm(x1)
    L.cgb200_invalidate(C.byref(A))


def test_tracked_policy_sees_every_edit_without_reading_unchanged_arrays(pkg, gpu_lib):
    # the default policy: the kernel tells whether the caller wrote to the arrays (csrc/writewatch.cpp)
    L = gpu_lib
    assert L.cgb200_set_matrix_cache(pkg.CACHE_TRACKED) == 0
    s = H.lap(420, 400)                       # values 6.7 MB, indices 3.4 MB: watched; row_ptr 0.7 MB: hashed
    A, _ = _structs(pkg, s)
    v = H.rhs(s.n, 33)
    vv = pkg.MvSparse(s.n, s.n, H.dptr(v), None, None)
    out = pkg.MvSparseP()

    def product_is_current():
        assert L.mv_mult(C.byref(A), C.byref(vv), C.byref(out)) == 0, pkg.last_error()
        assert np.array_equal(_bits(_take(out, s.n)), _bits(_spmv(s, v)))

    st0 = pkg.cache_stats()
    product_is_current()                      # upload (+ one hash, taken while the upload runs)
    st1 = pkg.cache_stats()
    assert st1["uploads"] == st0["uploads"] + 1
    for _ in range(3):
        product_is_current()                  # unchanged: settled by the kernel's answer alone
    st2 = pkg.cache_stats()
    assert st2["uploads"] == st1["uploads"]
    if st2["tracking_available"]:
        assert st2["tracked_clean"] == st1["tracked_clean"] + 3 and st2["hashes"] == st1["hashes"]
    else:
        assert st2["hashes"] == st1["hashes"] + 3          # no kernel support: the full hash, every call
    # edits: deep inside values, in the first bytes (a partial page: hashed, not watched), in the last
    # entry, in a column index, in row_ptr -- each is seen, without cgb200_invalidate
    k_mid = _off_grid_index(s.nnz)
    edits = [("values mid", lambda: s.vals.__setitem__(k_mid, 17.5)),
             ("values first", lambda: s.vals.__setitem__(1, -3.25)),
             ("values last", lambda: s.vals.__setitem__(s.nnz - 2, 0.125)),
             ("col index", lambda: s.cols.__setitem__(s.row_ptr[7] + 1, s.cols[s.row_ptr[7] + 1] + 1)),
             ("same values rewritten", lambda: s.vals.__setitem__(slice(1000, 200000), s.vals[1000:200000].copy()))]
    for name, edit in edits:
        before = pkg.cache_stats()
        edit()
        product_is_current()
        after = pkg.cache_stats()
        if name == "same values rewritten":
            assert after["uploads"] == before["uploads"], name      # written, hashed, found equal: no upload
            if after["tracking_available"]:
                assert after["hashes"] == before["hashes"] + 1 and after["tracked_dirty"] == before["tracked_dirty"] + 1
        else:
            assert after["uploads"] == before["uploads"] + 1, name
        product_is_current()                  # and the call after that is clean again
        again = pkg.cache_stats()
        assert again["uploads"] == after["uploads"]
        if again["tracking_available"]:
            assert again["tracked_clean"] == after["tracked_clean"] + 1 and again["hashes"] == after["hashes"], name
    # conj_grad goes through the same cache
    A2, b2 = _structs(pkg, s)
    xp = pkg.MvSparseP()
    assert L.conj_grad(10, C.byref(A2), C.byref(b2), C.byref(xp)) == 0
    x1 = _take(xp, s.n)
    s.vals[s.row_ptr[1:] - 1] *= 1.0
    d = np.flatnonzero(s.cols == np.repeat(np.arange(s.n), np.diff(s.row_ptr)))
    s.vals[d] = 4.75                           # the diagonal, in place
    assert L.conj_grad(10, C.byref(A2), C.byref(b2), C.byref(xp)) == 0
    x2 = _take(xp, s.n)
    xo = H.oracle_cg(s, 10)[0]
    assert np.linalg.norm(x2 - xo) <= TOL_X * np.linalg.norm(xo) and np.linalg.norm(x2 - x1) > 1e-3 * np.linalg.norm(x1)
    L.cgb200_invalidate(C.byref(A))
    L.cgb200_invalidate(C.byref(A2))
    L.cgb200_free_all(out)


def test_tracked_policy_with_two_structs_over_the_same_arrays(pkg, gpu_lib):
    # re-validating one struct clears the kernel's "written" marks of the shared arrays; the other
    # struct must not take the cleared marks for "unchanged" (abi.cu: WatchedRange.epoch)
    L = gpu_lib
    assert L.cgb200_set_matrix_cache(pkg.CACHE_TRACKED) == 0
    s = H.lap(400, 380)
    A1, _ = _structs(pkg, s)
    A2, _ = _structs(pkg, s)
    v = H.rhs(s.n, 4)
    vv = pkg.MvSparse(s.n, s.n, H.dptr(v), None, None)
    out = pkg.MvSparseP()

    def check(A):
        assert L.mv_mult(C.byref(A), C.byref(vv), C.byref(out)) == 0, pkg.last_error()
        assert np.array_equal(_bits(_take(out, s.n)), _bits(_spmv(s, v)))

    check(A1)
    check(A2)
    k = _off_grid_index(s.nnz)
    for first, second, val in ((A2, A1, 3.5), (A1, A2, -8.25), (A2, A1, 0.75)):
        s.vals[k] = val
        check(first)                           # sees the write, re-uploads, re-arms the shared range
        check(second)                          # the marks are gone; the epoch says: look again
        check(first)
        check(second)
    st0 = pkg.cache_stats()
    check(A1)
    check(A2)
    st1 = pkg.cache_stats()
    assert st1["uploads"] == st0["uploads"]
    if st1["tracking_available"]:
        assert st1["tracked_clean"] == st0["tracked_clean"] + 2 and st1["hashes"] == st0["hashes"]
    L.cgb200_invalidate(C.byref(A1))
    check(A2)                                  # releasing one struct leaves the other's watch intact
    s.vals[k + 1] = 9.0
    check(A2)
    L.cgb200_invalidate(C.byref(A2))
    L.cgb200_free_all(out)


def test_large_pageable_arrays_upload_through_the_staged_copy(pkg, gpu_lib, monkeypatch):
    # arrays of 64 MiB and more in ordinary memory go to the device in 4 MiB pieces from several host
    # threads (solver.cu: upload_host); the device copy is the same as the driver's own single copy
    L = gpu_lib
    s = H.lap(112, 110, 110)                  # 9.4 M entries: values 75 MB (staged), indices 38 MB (plain)
    assert 8 * s.nnz >= 64 << 20
    rng = np.random.default_rng(5)
    s.vals[:] = rng.random(s.nnz) - 0.5       # every piece distinguishable
    v = H.rhs(s.n, 12)
    want = _bits(_spmv(s, v))
    vv = pkg.MvSparse(s.n, s.n, H.dptr(v), None, None)
    for staged in ("1", "0"):
        monkeypatch.setenv("CGB200_STAGED_UPLOAD", staged)
        A, _ = _structs(pkg, s)
        out = pkg.MvSparseP()
        assert L.mv_mult(C.byref(A), C.byref(vv), C.byref(out)) == 0, pkg.last_error()
        assert np.array_equal(_bits(_take(out, s.n)), want), staged
        L.cgb200_invalidate(C.byref(A))
        L.cgb200_free_all(out)
    # arrays this large are usually backed by transparent huge pages: single stores anywhere in them are
    # still seen by the default (tracked) policy -- at the start, in the middle, at the end of a 2 MiB page
    monkeypatch.setenv("CGB200_STAGED_UPLOAD", "1")
    assert L.cgb200_set_matrix_cache(pkg.CACHE_TRACKED) == 0
    fresh = np.empty(s.nnz)                          # ask for huge pages whatever the system default is
    lo = (fresh.ctypes.data + 4095) & ~4095
    C.CDLL(None, use_errno=True).madvise(C.c_void_p(lo), C.c_size_t((fresh.ctypes.data + fresh.nbytes - lo) & ~4095), 14)
    fresh[:] = s.vals                                # (14 = MADV_HUGEPAGE; first touch happens here)
    s.vals = fresh
    want = _bits(_spmv(s, v))
    A, _ = _structs(pkg, s)
    out = pkg.MvSparseP()
    assert L.mv_mult(C.byref(A), C.byref(vv), C.byref(out)) == 0, pkg.last_error()
    assert np.array_equal(_bits(_take(out, s.n)), want)
    base = s.vals.ctypes.data
    first_huge = (-base) % (2 << 20) // 8            # index of the first element on a 2 MiB boundary
    for k in (first_huge, first_huge + (1 << 18) - 1, first_huge + 5 * (1 << 18) + 12345, s.nnz // 2, s.nnz - 3):
        u0 = pkg.cache_stats()["uploads"]
        s.vals[k] += 1.0
        assert L.mv_mult(C.byref(A), C.byref(vv), C.byref(out)) == 0, pkg.last_error()
        assert np.array_equal(_bits(_take(out, s.n)), _bits(_spmv(s, v))), k
        assert pkg.cache_stats()["uploads"] == u0 + 1
        assert L.mv_mult(C.byref(A), C.byref(vv), C.byref(out)) == 0
        assert pkg.cache_stats()["uploads"] == u0 + 1   # and clean again
    c0 = s.cols[s.nnz // 3]
    s.cols[s.nnz // 3] = c0 + 1 if c0 + 1 < s.n else c0 - 1
    assert L.mv_mult(C.byref(A), C.byref(vv), C.byref(out)) == 0, pkg.last_error()
    assert np.array_equal(_bits(_take(out, s.n)), _bits(_spmv(s, v)))
    L.cgb200_invalidate(C.byref(A))
    L.cgb200_free_all(out)


def test_large_vectors_move_through_the_staged_copy_in_both_directions(pkg, gpu_lib, monkeypatch):
    # b in and x out of 32 MiB and more (pageable) are staged like the matrix arrays; same results
    L = gpu_lib
    s = H.lap(2100, 2048)                     # 4.3 M unknowns: 34 MB per vector
    assert 8 * s.n >= 32 << 20
    v = H.rhs(s.n, 9)
    want = _bits(_spmv(s, v))
    xo = H.oracle_cg(s, 6)[0]
    vv = pkg.MvSparse(s.n, s.n, H.dptr(v), None, None)
    got = {}
    for staged in ("1", "0"):
        monkeypatch.setenv("CGB200_STAGED_UPLOAD", staged)
        A, b = _structs(pkg, s)
        out, xp = pkg.MvSparseP(), pkg.MvSparseP()
        assert L.mv_mult(C.byref(A), C.byref(vv), C.byref(out)) == 0, pkg.last_error()
        assert np.array_equal(_bits(_take(out, s.n)), want), staged
        assert L.conj_grad(6, C.byref(A), C.byref(b), C.byref(xp)) == 0, pkg.last_error()
        got[staged] = _take(xp, s.n)
        assert np.linalg.norm(got[staged] - xo) <= TOL_X * np.linalg.norm(xo)
        L.cgb200_invalidate(C.byref(A))
        L.cgb200_free_all(out)
        L.cgb200_free_all(xp)
    assert np.array_equal(_bits(got["1"]), _bits(got["0"]))


def test_pinned_and_sampled_matrices_follow_the_documented_contract(pkg, strict):
    L = strict
    s = H.lap(40, 30)
    A, _ = _structs(pkg, s)
    v = H.rhs(s.n, 32)
    vv = pkg.MvSparse(s.n, s.n, H.dptr(v), None, None)
    out = pkg.MvSparseP()
    want_old = _spmv(s, v)
    k = _off_grid_index(s.nnz)
    # cgb200_to_device(A): the caller declares the device copy current until cgb200_invalidate
    assert L.cgb200_to_device(C.byref(A)) == 0, pkg.last_error()
    s.vals[k] = -77.0
    assert L.mv_mult(C.byref(A), C.byref(vv), C.byref(out)) == 0
    assert np.array_equal(_bits(_take(out, s.n)), _bits(want_old))        # pinned: the device copy answers
    assert L.cgb200_invalidate(C.byref(A)) == 0
    assert L.mv_mult(C.byref(A), C.byref(vv), C.byref(out)) == 0
    want_new = _spmv(s, v)
    assert np.array_equal(_bits(_take(out, s.n)), _bits(want_new))
    # sampled policy: a rebuilt matrix (sampled entries change) is noticed without invalidate
    assert L.cgb200_set_matrix_cache(pkg.CACHE_SAMPLED) == 0
    s.vals[:] = s.vals * 2.0
    assert L.mv_mult(C.byref(A), C.byref(vv), C.byref(out)) == 0
    assert np.array_equal(_bits(_take(out, s.n)), _bits(_spmv(s, v)))
    # off: every call uploads
    assert L.cgb200_set_matrix_cache(pkg.CACHE_OFF) == 0
    s.vals[k] = 5.0
    assert L.mv_mult(C.byref(A), C.byref(vv), C.byref(out)) == 0
    assert np.array_equal(_bits(_take(out, s.n)), _bits(_spmv(s, v)))
    assert L.cgb200_set_matrix_cache(7) == -1
    L.cgb200_invalidate(C.byref(A))
    L.cgb200_free_all(out)


@pytest.mark.parametrize("mk,M", [(lambda: H.lap(20, 20), 40), (lambda: H.lap(9, 8, 7), 25), (lambda: H.banded(400, 70), 6),
                                  (lambda: H.dense(256), 3)])
def test_negative_definite_system_solves_like_the_reference(pkg, gpu_lib, mk, M):
    # (-A, -b): the reference's loop divides by a negative p.Ap and returns the x of (A, b)
    L = gpu_lib
    neg = H.negated(mk())
    A, b = _structs(pkg, neg)
    xp = pkg.MvSparseP()
    o = pkg.Opts(M, 0.0, 0, 0, 0)
    r = pkg.Result()
    assert L.cgb200_solve(C.byref(A), C.byref(b), C.byref(xp), C.byref(o), C.byref(r)) == 0, pkg.last_error()
    x = _take(xp, neg.n)
    L.cgb200_invalidate(C.byref(A))
    assert r.status == pkg.MAXITER and r.iterations == M + 1
    want = H.ref_cg(neg, M) if H.have_ref() else H.oracle_cg(neg, M)[0]
    assert np.all(np.isfinite(want))
    assert np.linalg.norm(x - want) <= TOL_X * np.linalg.norm(want)
    # and with a tolerance, in the single-reduction recurrence too
    for variant in (0, 1):
        o = pkg.Opts(2000, 1e-10, variant, 0, 0)
        assert L.cgb200_solve(C.byref(A), C.byref(b), C.byref(xp), C.byref(o), C.byref(r)) == 0, pkg.last_error()
        assert r.status == pkg.CONVERGED
        assert H.true_relres(neg, _take(xp, neg.n)) <= 1.05e-10
        L.cgb200_invalidate(C.byref(A))


def test_malformed_csr_is_refused_and_the_library_stays_usable(pkg, gpu_lib):
    L = gpu_lib
    s = H.lap(12, 12)
    v = H.rhs(s.n, 3)
    vv = pkg.MvSparse(s.n, s.n, H.dptr(v), None, None)
    out = pkg.MvSparseP()
    one_based = H.System(s.row_ptr, s.cols + 1, s.vals, s.b)             # a 1-based file: column n
    A, b = _structs(pkg, one_based)
    assert L.mv_mult(C.byref(A), C.byref(vv), C.byref(out)) == -1
    assert "malformed CSR" in pkg.last_error()
    negc = H.System(s.row_ptr, s.cols - 1, s.vals, s.b)                   # column -1
    A2, _ = _structs(pkg, negc)
    assert L.conj_grad(3, C.byref(A2), C.byref(b), C.byref(out)) == -1
    rp = s.row_ptr.copy()
    rp[5], rp[6] = rp[6], rp[5] - 1                                       # a decreasing step
    bad_rp = H.System(rp, s.cols, s.vals, s.b)
    A3, _ = _structs(pkg, bad_rp)
    assert L.mv_mult(C.byref(A3), C.byref(vv), C.byref(out)) == -1
    assert "row_ptr" in pkg.last_error()
    with pytest.raises(pkg.CgError):
        pkg.DeviceSystem.from_csr(s.n, 0, s.row_ptr, s.cols + 1, s.vals, s.b)
    # nothing is broken: the well-formed matrix runs, bit-exact
    A4, _ = _structs(pkg, s)
    assert L.mv_mult(C.byref(A4), C.byref(vv), C.byref(out)) == 0, pkg.last_error()
    assert np.array_equal(_bits(_take(out, s.n)), _bits(_spmv(s, v)))
    L.cgb200_invalidate(C.byref(A4))


@pytest.mark.parametrize("M", [0, 3, 31, 32, 40, 70])
def test_fixed_count_solve_launches_exactly_the_iterations_it_runs(pkg, gpu_lib, M):
    # max_iter = M is M + 1 updates (cg.c:110-135): 3 kernels each + the loop entry, no early-exit
    # launches behind the last one (the final batch is a tail graph of the right length); a system this
    # small runs its two vector kernels as one launch by default (k_update_xrp): 2 per update
    try:
        for fused, per in ((0, 3), (1, 2)):
            gpu_lib.cgb200_set_fused_vector_update(fused)
            d = pkg.DeviceSystem.lap(64, 48)
            r = d.solve(M)
            assert r.iterations == M + 1 and r.status == pkg.MAXITER
            assert r.kernel_launches == per * (M + 1) + 1
            t = d.solve_async(M)
            ra = d.solve_wait(t)
            assert ra.iterations == M + 1 and ra.kernel_launches == per * (M + 1) + 1
            d.free()
    finally:
        gpu_lib.cgb200_set_fused_vector_update(1)


def test_calls_from_several_threads_are_serialised_not_corrupted(pkg, gpu_lib):
    # the reference's operators are re-entrant (mv_ops.c has no globals); here every entry point
    # takes the process-wide lock (include/cgb200.h: Threading)
    L = gpu_lib
    systems = [H.lap(30 + 3 * i, 20 + i) for i in range(4)]
    want = [H.oracle_cg(s, 25)[0] for s in systems]
    devs = [pkg.DeviceSystem.from_csr(s.n, 0, s.row_ptr, s.cols, s.vals, s.b) for s in systems]
    errs = []

    def work(i):
        try:
            for _ in range(6):
                r = devs[i].solve(25)
                x = devs[i].x()
                assert r.iterations == 26
                assert np.linalg.norm(x - want[i]) <= TOL_X * np.linalg.norm(want[i])
                a = H.rhs(systems[i].n, 40 + i)
                va = pkg.MvSparse(len(a), len(a), H.dptr(a), None, None)
                got = L.dot_product(C.byref(va), C.byref(va))
                assert abs(got - float(a @ a)) <= 1e-12 * float(a @ a)
        except Exception as e:  # noqa: BLE001
            errs.append((i, repr(e)))

    th = [threading.Thread(target=work, args=(i,)) for i in range(4)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    for d in devs:
        d.free()
    assert not errs, errs
