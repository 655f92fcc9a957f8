"""Shared test plumbing: ctypes views of the oracle, the compiled reference and the product.

Test infrastructure.  The oracle (oracle/liboracle.so) and the compiled, unmodified reference
(oracle/_ref/libcgref.so, built from /root/reference by oracle/Makefile) are checkers only.
"""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
ORACLE_SO = os.path.join(ORACLE_DIR, "liboracle.so")
REF_SO = os.path.join(ORACLE_DIR, "_ref", "libcgref.so")
REF_CG = os.path.join(ORACLE_DIR, "_ref", "cg")
GOLDEN = os.path.join(ROOT, "tests", "golden")

c_double_p = C.POINTER(C.c_double)
c_int_p = C.POINTER(C.c_int)


class MvSparse(C.Structure):
    """struct __mv_sparse (reference mv_ops.h:17-23): the ABI type of the path."""

    _fields_ = [
        ("size", C.c_int),
        ("nnz", C.c_int),
        ("values", c_double_p),
        ("col_indices", c_int_p),
        ("row_ptr", c_int_p),
    ]


MvSparseP = C.POINTER(MvSparse)


def dptr(a):
    return a.ctypes.data_as(c_double_p)


def iptr(a):
    return a.ctypes.data_as(c_int_p)


def build_oracle():
    if not os.path.exists(ORACLE_SO) or os.path.getmtime(ORACLE_SO) < os.path.getmtime(
        os.path.join(ORACLE_DIR, "cg_oracle.c")
    ):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "-s"], stdout=subprocess.DEVNULL)


_oracle = None


def oracle():
    """ctypes handle on oracle/liboracle.so with argtypes declared."""
    global _oracle
    if _oracle is not None:
        return _oracle
    build_oracle()
    L = C.CDLL(ORACLE_SO)
    L.oracle_uniform.restype = C.c_double
    L.oracle_uniform.argtypes = [C.c_uint64, C.c_uint64]
    L.oracle_gen_rhs.argtypes = [C.c_int64, C.c_int, C.c_uint64, c_double_p]
    L.oracle_lap_nnz.restype = C.c_int64
    L.oracle_lap_nnz.argtypes = [C.c_int, C.c_int, C.c_int]
    L.oracle_gen_lap.argtypes = [C.c_int, C.c_int, C.c_int, c_int_p, c_int_p, c_double_p]
    L.oracle_gen_dense.argtypes = [C.c_int, C.c_uint64, c_int_p, c_int_p, c_double_p]
    L.oracle_gen_banded.restype = C.c_int64
    L.oracle_gen_banded.argtypes = [C.c_int, C.c_int, C.c_uint64, c_int_p, c_int_p, c_double_p]
    L.oracle_spmv.argtypes = [C.c_int, c_int_p, c_int_p, c_double_p, c_double_p, c_double_p]
    L.oracle_dot.restype = C.c_double
    L.oracle_dot.argtypes = [C.c_int, c_double_p, c_double_p]
    L.oracle_scale.argtypes = [C.c_int, C.c_double, c_double_p, c_double_p]
    L.oracle_add.argtypes = [C.c_int, c_double_p, c_double_p, c_double_p]
    L.oracle_sub.argtypes = [C.c_int, c_double_p, c_double_p, c_double_p]
    L.oracle_conj_grad.restype = C.c_int
    L.oracle_conj_grad.argtypes = [
        C.c_int, c_int_p, c_int_p, c_double_p, c_double_p, C.c_int, C.c_double, C.c_int,
        c_double_p, c_double_p, c_int_p,
    ]
    L.oracle_conj_grad_sr.restype = C.c_int
    L.oracle_conj_grad_sr.argtypes = [
        C.c_int, c_int_p, c_int_p, c_double_p, c_double_p, C.c_int, C.c_double,
        c_double_p, c_double_p, c_int_p,
    ]
    L.oracle_pcg.restype = C.c_int
    L.oracle_pcg.argtypes = [
        C.c_int, c_int_p, c_int_p, c_double_p, c_double_p, c_double_p, c_double_p, C.c_int, C.c_double,
        c_double_p, c_double_p, c_int_p,
    ]
    L.oracle_diag_inv.argtypes = [C.c_int, c_int_p, c_int_p, c_double_p, c_double_p]
    L.oracle_true_relres.restype = C.c_double
    L.oracle_true_relres.argtypes = [C.c_int, c_int_p, c_int_p, c_double_p, c_double_p, c_double_p]
    L.oracle_num_threads.restype = C.c_int
    L.oracle_cg_iterations_mt.restype = C.c_double
    L.oracle_cg_iterations_mt.argtypes = [
        C.c_int, c_int_p, c_int_p, c_double_p, c_double_p, C.c_int, c_double_p,
    ]
    L.oracle_cg_iterations_dense_mt.restype = C.c_double
    L.oracle_cg_iterations_dense_mt.argtypes = [C.c_int, c_double_p, c_double_p, C.c_int, c_double_p]
    _oracle = L
    return L


def have_ref():
    return os.path.exists(REF_SO)


_ref = None


def ref():
    """ctypes handle on the compiled, unmodified reference (oracle/_ref/libcgref.so)."""
    global _ref
    if _ref is not None:
        return _ref
    L = C.CDLL(REF_SO, mode=C.RTLD_LOCAL)
    L.new_mv_struct.restype = MvSparseP
    L.new_mv_struct_with_size.restype = MvSparseP
    L.new_mv_struct_with_size.argtypes = [C.c_int]
    L.mv_deep_copy.restype = MvSparseP
    L.mv_deep_copy.argtypes = [MvSparseP]
    L.mat_get_row.argtypes = [MvSparseP, C.c_int, c_double_p]
    L.dot_product.restype = C.c_double
    L.dot_product.argtypes = [MvSparseP, MvSparseP]
    L.sv_mult.argtypes = [C.c_double, MvSparseP, C.POINTER(MvSparseP)]
    L.mv_mult.argtypes = [MvSparseP, MvSparseP, C.POINTER(MvSparseP)]
    L.vec_add.argtypes = [MvSparseP, MvSparseP, C.POINTER(MvSparseP)]
    L.vec_sub.argtypes = [MvSparseP, MvSparseP, C.POINTER(MvSparseP)]
    L.conj_grad.argtypes = [C.c_int, MvSparseP, MvSparseP, C.POINTER(MvSparseP)]
    L.read_input_file.argtypes = [C.c_char_p, MvSparseP, MvSparseP]
    _ref = L
    return L


# ---------------------------------------------------------------------------------------------
# Systems
# ---------------------------------------------------------------------------------------------


class System:
    """A CSR system held in numpy arrays (kept alive while ctypes structs point into them)."""

    def __init__(self, row_ptr, cols, vals, b, name=""):
        self.row_ptr = np.ascontiguousarray(row_ptr, dtype=np.int32)
        self.cols = np.ascontiguousarray(cols, dtype=np.int32)
        self.vals = np.ascontiguousarray(vals, dtype=np.float64)
        self.b = np.ascontiguousarray(b, dtype=np.float64)
        self.n = len(self.row_ptr) - 1
        self.nnz = len(self.vals)
        self.name = name

    def A_struct(self):
        return MvSparse(self.n, self.nnz, dptr(self.vals), iptr(self.cols), iptr(self.row_ptr))

    def b_struct(self):
        return MvSparse(self.n, self.n, dptr(self.b), None, None)


def rhs(n, seed=7, row0=0):
    b = np.empty(n, dtype=np.float64)
    oracle().oracle_gen_rhs(row0, n, seed, dptr(b))
    return b


def lap(nx, ny, nz=1, seed=7):
    O = oracle()
    n = nx * ny * nz
    nnz = O.oracle_lap_nnz(nx, ny, nz)
    rp = np.empty(n + 1, dtype=np.int32)
    ci = np.empty(nnz, dtype=np.int32)
    va = np.empty(nnz, dtype=np.float64)
    O.oracle_gen_lap(nx, ny, nz, iptr(rp), iptr(ci), dptr(va))
    return System(rp, ci, va, rhs(n, seed), name=f"lap_{nx}x{ny}x{nz}")


def dense(n, seed=1, bseed=7):
    O = oracle()
    rp = np.empty(n + 1, dtype=np.int32)
    ci = np.empty(n * n, dtype=np.int32)
    va = np.empty(n * n, dtype=np.float64)
    O.oracle_gen_dense(n, seed, iptr(rp), iptr(ci), dptr(va))
    return System(rp, ci, va, rhs(n, bseed), name=f"dense_{n}")


def banded(n, half_bw, seed=3, bseed=7):
    """Symmetric banded SPD, strictly diagonally dominant, irreducible (bleed-free for the
    reference's mat_get_row): off-diagonals U(-1/2,1/2) by (min,max) pair, diagonal = bw + 1."""
    O = oracle()
    nnz = O.oracle_gen_banded(n, half_bw, seed, None, None, None)
    rp = np.empty(n + 1, dtype=np.int32)
    ci = np.empty(nnz, dtype=np.int32)
    va = np.empty(nnz, dtype=np.float64)
    O.oracle_gen_banded(n, half_bw, seed, iptr(rp), iptr(ci), dptr(va))
    return System(rp, ci, va, rhs(n, bseed), name=f"banded_{n}_{half_bw}")


def toeplitz(n, coeffs, bseed=7):
    """Symmetric Toeplitz band: A[i, i+k] = A[i+k, i] = coeffs[k] (coeffs[0] the diagonal), clipped at
    the edges.  Few distinct rows once written as (column - row, value): a pattern-format case."""
    h = len(coeffs) - 1
    rp, ci, va = [0], [], []
    for i in range(n):
        for j in range(max(0, i - h), min(n, i + h + 1)):
            ci.append(j)
            va.append(coeffs[abs(j - i)])
        rp.append(len(ci))
    return System(rp, ci, va, rhs(n, bseed), name=f"toeplitz_{n}_{h}")


def negated(sysm):
    """(-A, -b): a negative-definite system with the same solution.  The reference divides by p.Ap
    whatever its sign (cg.c:113), so its iterates are those of (A, b) with r, p, s negated."""
    return System(sysm.row_ptr, sysm.cols, -sysm.vals, -sysm.b, name="neg_" + sysm.name)


def tridiag4():
    """The 4x4 known-answer system: tridiag(-1,2,-1), b = 1."""
    rp = [0, 2, 5, 8, 10]
    ci = [0, 1, 0, 1, 2, 1, 2, 3, 2, 3]
    va = [2, -1, -1, 2, -1, -1, 2, -1, -1, 2]
    return System(rp, ci, va, np.ones(4), name="tridiag4")


# ---------------------------------------------------------------------------------------------
# Running the checkers
# ---------------------------------------------------------------------------------------------


def oracle_cg(sysm, max_iter, tol=0.0, guard=0):
    """Returns (x, updates, rr, status) from the restatement (oracle/cg_oracle.c)."""
    x = np.empty(sysm.n, dtype=np.float64)
    rr = C.c_double(0.0)
    st = C.c_int(0)
    it = oracle().oracle_conj_grad(
        sysm.n, iptr(sysm.row_ptr), iptr(sysm.cols), dptr(sysm.vals), dptr(sysm.b),
        max_iter, tol, guard, dptr(x), C.byref(rr), C.byref(st),
    )
    return x, it, rr.value, st.value


def diag_inv(sysm):
    d = np.empty(sysm.n, dtype=np.float64)
    oracle().oracle_diag_inv(sysm.n, iptr(sysm.row_ptr), iptr(sysm.cols), dptr(sysm.vals), dptr(d))
    return d


def oracle_pcg(sysm, max_iter, tol=0.0, x0=None, jacobi=False):
    """Returns (x, updates, rr, status) from the preconditioned / warm-started restatement."""
    x = np.empty(sysm.n, dtype=np.float64)
    rr = C.c_double(0.0)
    st = C.c_int(0)
    dinv = diag_inv(sysm) if jacobi else None
    x0c = np.ascontiguousarray(x0, dtype=np.float64) if x0 is not None else None
    it = oracle().oracle_pcg(
        sysm.n, iptr(sysm.row_ptr), iptr(sysm.cols), dptr(sysm.vals), dptr(sysm.b),
        dptr(x0c) if x0c is not None else None, dptr(dinv) if dinv is not None else None,
        max_iter, tol, dptr(x), C.byref(rr), C.byref(st))
    return x, it, rr.value, st.value


def oracle_cg_sr(sysm, max_iter, tol=0.0):
    """Same as oracle_cg for the single-reduction recurrence (not in the reference; restates the
    product's optional variant so that it can be compared with something that rounds alike)."""
    x = np.empty(sysm.n, dtype=np.float64)
    rr = C.c_double(0.0)
    st = C.c_int(0)
    it = oracle().oracle_conj_grad_sr(
        sysm.n, iptr(sysm.row_ptr), iptr(sysm.cols), dptr(sysm.vals), dptr(sysm.b),
        max_iter, tol, dptr(x), C.byref(rr), C.byref(st),
    )
    return x, it, rr.value, st.value


def ref_cg(sysm, max_iter):
    """x from the unmodified reference's conj_grad (cg.c:88), read from memory (not stdout)."""
    R = ref()
    A, b = sysm.A_struct(), sysm.b_struct()
    xp = MvSparseP()
    rc = R.conj_grad(max_iter, C.byref(A), C.byref(b), C.byref(xp))
    assert rc == 0
    return np.ctypeslib.as_array(xp.contents.values, shape=(sysm.n,)).copy()


def true_relres(sysm, x):
    x = np.ascontiguousarray(x, dtype=np.float64)
    return oracle().oracle_true_relres(
        sysm.n, iptr(sysm.row_ptr), iptr(sysm.cols), dptr(sysm.vals), dptr(sysm.b), dptr(x)
    )


def write_text_format(sysm, path):
    """The reference's 4-line input format (cg.c:173-201): cols / row_ptr / values / b, comma
    separated, every line (including the 4th) newline-terminated, tokens < 64 chars."""
    with open(path, "w") as f:
        f.write(",".join(str(int(v)) for v in sysm.cols) + "\n")
        f.write(",".join(str(int(v)) for v in sysm.row_ptr) + "\n")
        f.write(",".join(repr(float(v)) for v in sysm.vals) + "\n")
        f.write(",".join(repr(float(v)) for v in sysm.b) + "\n")
