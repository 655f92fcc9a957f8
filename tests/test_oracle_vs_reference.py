"""CPU: the oracle against the compiled, unmodified reference (oracle/_ref/libcgref.so), live.
Runs wherever oracle/_ref exists (built here from /root/reference; the prebuilt files travel to the
GPU box).  This is the "restatement self-check" tier of SURVEY.md section 4."""
import ctypes as C

import numpy as np
import pytest

import helpers as H

pytestmark = pytest.mark.skipif(not H.have_ref(), reason="oracle/_ref not built (no reference tree here)")


def _bits(a):
    return np.ascontiguousarray(a, dtype=np.float64).view(np.uint64)


CASES = [
    (lambda: H.lap(16, 16), 40),
    (lambda: H.lap(48, 20), 25),
    (lambda: H.lap(10, 10, 10), 30),
    (lambda: H.banded(500, 3), 10),
    (lambda: H.banded(200, 60), 6),
    (lambda: H.dense(300), 3),
    (lambda: H.lap(5, 1), 2),
]


@pytest.mark.parametrize("mk,M", CASES)
def test_loop_bitwise(mk, M):
    s = mk()
    assert np.array_equal(_bits(H.ref_cg(s, M)), _bits(H.oracle_cg(s, M)[0]))


@pytest.mark.parametrize("mk,M", [(lambda: H.lap(16, 16), 40), (lambda: H.lap(6, 5, 4), 12), (lambda: H.banded(200, 9), 8)])
def test_negative_definite_system_is_not_a_breakdown(mk, M):
    # the reference divides by p.Ap regardless of its sign (cg.c:113): on (-A, -b) it walks the
    # iterates of (A, b).  The product's breakdown guard (oracle guard = 1 restates it) must let a
    # negative p.Ap through and stop only where the division has no finite result.
    s = mk()
    neg = H.negated(s)
    x_ref = H.ref_cg(neg, M)
    x_g, it, _, st = H.oracle_cg(neg, M, guard=1)
    assert st == 0 and it == M + 1
    assert np.array_equal(_bits(x_ref), _bits(x_g))
    assert np.array_equal(_bits(x_ref), _bits(H.ref_cg(s, M)))  # same x, bit for bit: signs cancel


def test_reference_runs_its_own_operators():
    # libcgb200.so exports the same symbol names by design; the reference library must not bind to
    # them (oracle/Makefile links it -Bsymbolic).  A CPU-only box proves it: the product's
    # operators fail without a device, the reference's do not.
    R = H.ref()
    a = H.rhs(10, 1)
    va = H.MvSparse(10, 10, H.dptr(a), None, None)
    assert R.dot_product(C.byref(va), C.byref(va)) == H.oracle().oracle_dot(10, H.dptr(a), H.dptr(a))


def test_reference_text_reader_round_trip(tmp_path):
    # the 4-line input format (cg.c:173-201) as written by helpers.write_text_format
    s = H.lap(6, 5)
    p = tmp_path / "in.txt"
    H.write_text_format(s, str(p))
    R = H.ref()
    A = R.new_mv_struct()
    b = R.new_mv_struct()
    assert R.read_input_file(str(p).encode(), A, b) == 0
    assert A.contents.size == s.n and A.contents.nnz == s.nnz and b.contents.size == s.n
    assert np.array_equal(np.ctypeslib.as_array(A.contents.values, shape=(s.nnz,)), s.vals)
    assert np.array_equal(np.ctypeslib.as_array(A.contents.col_indices, shape=(s.nnz,)), s.cols)
    assert np.array_equal(np.ctypeslib.as_array(A.contents.row_ptr, shape=(s.n + 1,)), s.row_ptr)
    assert np.array_equal(np.ctypeslib.as_array(b.contents.values, shape=(s.n,)), s.b)
