"""conjugate-gradient_b200: host-side mirror of the reference's operator interface over libcgb200.so.

The product is the C-ABI shared library built from csrc/ (hand-written sm_100a CUDA kernels behind
the reference's mv_ops.h API, include/mv_ops.h + include/cgb200.h).  This module is the thin
ctypes view of that library used by the tests and bench.py; names, argument meaning and error
behaviour are the reference's (mv_ops.h:26-42, cg.c:24).  Nothing here computes: if the library is
missing or no CUDA device is usable, calls fail loudly.

The directory name carries a hyphen, so import it through `__graft_entry__.load_package()` (module
name `conjugate_gradient_b200`).
"""
import ctypes as C
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
# CGB200_LIBRARY: another build of the same library (A/B comparisons of kernel changes, tools/ab_build.sh)
LIB_PATH = os.environ.get("CGB200_LIBRARY") or os.path.join(HERE, "libcgb200.so")
CG_DRIVER = os.path.join(HERE, "cg")

c_double_p = C.POINTER(C.c_double)
c_int_p = C.POINTER(C.c_int)


class MvSparse(C.Structure):
    """struct __mv_sparse (include/mv_ops.h; reference mv_ops.h:17-23)."""

    _fields_ = [
        ("size", C.c_int),
        ("nnz", C.c_int),
        ("values", c_double_p),
        ("col_indices", c_int_p),
        ("row_ptr", c_int_p),
    ]


MvSparseP = C.POINTER(MvSparse)


class Opts(C.Structure):
    """cgb200_opts (include/cgb200.h)."""

    _fields_ = [
        ("max_iter", C.c_int),
        ("tol", C.c_double),
        ("variant", C.c_int),
        ("check_every", C.c_int),
        ("flags", C.c_int),
    ]


class Result(C.Structure):
    """cgb200_result (include/cgb200.h)."""

    _fields_ = [
        ("iterations", C.c_int),
        ("status", C.c_int),
        ("rr", C.c_double),
        ("bb", C.c_double),
        ("relres", C.c_double),
        ("seconds", C.c_double),
        ("bytes_per_iter", C.c_double),
        ("path", C.c_int),
        ("world", C.c_int),
        ("kernel_launches", C.c_longlong),
    ]


RUNNING, MAXITER, CONVERGED, BREAKDOWN, EXACT, COMM_TIMEOUT = 0, 1, 2, 3, 4, 5
CLASSIC, SINGLE_REDUCTION = 0, 1
EAGER, RESIDENT = 0, 1
OPT_JACOBI, OPT_X0 = 1, 2
CACHE_STRICT, CACHE_SAMPLED, CACHE_OFF, CACHE_TRACKED = 0, 1, 2, 3
DOT_STATIC_MAX_N = 1 << 28

# every symbol include/mv_ops.h and include/cgb200.h declare (checked by tests/test_abi_symbols.py)
MV_OPS_SYMBOLS = [
    "new_mv_struct", "new_mv_struct_with_size", "free_mv_struct", "mv_deep_copy", "print_sparse",
    "mat_get_row", "dot_product", "sv_mult", "mv_mult", "vec_add", "vec_sub", "conj_grad",
]


class LibraryMissing(RuntimeError):
    pass


def build(verbose=False):
    """Compile libcgb200.so (and the cg driver) in-tree for sm_100a with nvcc."""
    out = None if verbose else subprocess.DEVNULL
    subprocess.check_call(["make", "-C", HERE, "all"], stdout=out)


_lib = None


def lib():
    """The loaded libcgb200.so with argtypes declared.  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise LibraryMissing(
            f"{LIB_PATH} not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU fallback)"
        )
    L = C.CDLL(LIB_PATH, mode=C.RTLD_LOCAL)
    # --- reference surface ---
    L.new_mv_struct.restype = MvSparseP
    L.new_mv_struct.argtypes = []
    L.new_mv_struct_with_size.restype = MvSparseP
    L.new_mv_struct_with_size.argtypes = [C.c_int]
    L.free_mv_struct.restype = None
    L.free_mv_struct.argtypes = [MvSparseP]
    L.mv_deep_copy.restype = MvSparseP
    L.mv_deep_copy.argtypes = [MvSparseP]
    L.print_sparse.restype = None
    L.print_sparse.argtypes = [MvSparseP]
    L.mat_get_row.restype = C.c_int
    L.mat_get_row.argtypes = [MvSparseP, C.c_int, c_double_p]
    L.dot_product.restype = C.c_double
    L.dot_product.argtypes = [MvSparseP, MvSparseP]
    L.sv_mult.restype = C.c_int
    L.sv_mult.argtypes = [C.c_double, MvSparseP, C.POINTER(MvSparseP)]
    for name in ("mv_mult", "vec_add", "vec_sub"):
        f = getattr(L, name)
        f.restype = C.c_int
        f.argtypes = [MvSparseP, MvSparseP, C.POINTER(MvSparseP)]
    L.conj_grad.restype = C.c_int
    L.conj_grad.argtypes = [C.c_int, MvSparseP, MvSparseP, C.POINTER(MvSparseP)]
    # --- extensions ---
    L.cgb200_init.argtypes = [C.c_int]
    L.cgb200_shutdown.restype = None
    L.cgb200_last_error.restype = C.c_char_p
    L.cgb200_launch_count.restype = C.c_longlong
    L.cgb200_device_available.restype = C.c_int
    L.cgb200_read_system.argtypes = [C.c_char_p, MvSparseP, MvSparseP]
    L.cgb200_write_system.argtypes = [C.c_char_p, MvSparseP, MvSparseP, C.c_int]
    L.cgb200_guard_check.restype = C.c_longlong
    L.cgb200_guard_check.argtypes = []
    L.cgb200_guard_selftest.restype = C.c_int
    L.cgb200_guard_selftest.argtypes = []
    L.cgb200_dot_static.argtypes = [C.c_void_p, C.c_void_p, C.c_longlong, C.c_void_p, C.c_void_p]
    L.cgb200_set_gpus.argtypes = [C.c_int]
    L.cgb200_set_mode.argtypes = [C.c_int]
    L.cgb200_set_matrix_cache.argtypes = [C.c_int]
    L.cgb200_cache_stats.argtypes = [C.POINTER(C.c_longlong)]
    L.cgb200_write_tracking_op.argtypes = [C.c_void_p, C.c_size_t, C.c_int]
    L.cgb200_write_tracking_op.restype = C.c_int
    L.cgb200_content_hash.restype = C.c_ulonglong
    L.cgb200_content_hash.argtypes = [MvSparseP]
    for name in ("cgb200_to_device", "cgb200_to_host", "cgb200_invalidate"):
        getattr(L, name).argtypes = [MvSparseP]
    L.cgb200_free_all.restype = None
    L.cgb200_free_all.argtypes = [MvSparseP]
    L.cgb200_solve.argtypes = [MvSparseP, MvSparseP, C.POINTER(MvSparseP), C.POINTER(Opts), C.POINTER(Result)]
    L.cgb200_system_lap.restype = C.c_void_p
    L.cgb200_system_lap.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint64]
    L.cgb200_system_dense.restype = C.c_void_p
    L.cgb200_system_dense.argtypes = [C.c_int, C.c_int, C.c_int, C.c_uint64, C.c_uint64]
    L.cgb200_system_from_csr.restype = C.c_void_p
    L.cgb200_system_from_csr.argtypes = [C.c_longlong, C.c_longlong, C.c_int, c_int_p, c_int_p, c_double_p, c_double_p]
    L.cgb200_system_free.restype = None
    L.cgb200_system_free.argtypes = [C.c_void_p]
    L.cgb200_system_set_rhs.argtypes = [C.c_void_p, c_double_p]
    L.cgb200_system_gen_rhs.argtypes = [C.c_void_p, C.c_uint64]
    L.cgb200_system_get_x.argtypes = [C.c_void_p, c_double_p]
    L.cgb200_system_set_x.argtypes = [C.c_void_p, c_double_p]
    L.cgb200_system_solve.argtypes = [C.c_void_p, C.POINTER(Opts), C.POINTER(Result)]
    L.cgb200_system_solve_async.argtypes = [C.c_void_p, C.POINTER(Opts), C.POINTER(C.c_longlong)]
    L.cgb200_solve_wait.argtypes = [C.c_longlong, C.POINTER(Result)]
    L.cgb200_system_matvec.argtypes = [C.c_void_p, c_double_p, c_double_p]
    L.cgb200_system_true_relres.argtypes = [C.c_void_p, c_double_p]
    L.cgb200_system_time_kernel.argtypes = [C.c_void_p, C.c_int, C.c_int, c_double_p]
    L.cgb200_system_bytes.argtypes = [C.c_void_p, c_double_p, c_double_p]
    L.cgb200_system_comm_bytes.restype = C.c_double
    L.cgb200_system_comm_bytes.argtypes = [C.c_void_p]
    L.cgb200_system_trace.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_ulonglong)]
    L.cgb200_system_download.argtypes = [C.c_void_p, c_double_p, c_int_p, c_int_p]
    L.cgb200_system_get_b.argtypes = [C.c_void_p, c_double_p]
    L.cgb200_timer_stop.argtypes = [c_double_p]
    L.cgb200_system_rows.restype = C.c_longlong
    L.cgb200_system_rows.argtypes = [C.c_void_p]
    L.cgb200_system_nnz.restype = C.c_longlong
    L.cgb200_system_nnz.argtypes = [C.c_void_p]
    L.cgb200_plan_partition.argtypes = [C.c_int, C.c_int, C.c_longlong, C.POINTER(C.c_longlong), C.c_int, C.POINTER(C.c_longlong)]
    L.cgb200_plan_push.argtypes = [C.c_int, C.c_int, C.c_longlong, C.POINTER(C.c_longlong), C.c_int, C.POINTER(C.c_longlong)]
    L.cgb200_dist_unique_id.argtypes = [C.c_void_p]
    L.cgb200_dist_init.argtypes = [C.c_int, C.c_int, C.c_void_p]
    L.cgb200_dist_comm_mode.restype = C.c_char_p
    L.cgb200_system_comm_is_p2p.argtypes = [C.c_void_p]
    L.cgb200_system_one_kernel_iteration.argtypes = [C.c_void_p]
    L.cgb200_set_fused_vector_update.argtypes = [C.c_int]
    L.cgb200_set_fused_vector_update.restype = None
    L.cgb200_system_use_pattern_format.argtypes = [C.c_void_p, C.c_int]
    L.cgb200_system_use_banded_format.argtypes = [C.c_void_p, C.c_int]
    _lib = L
    return L


def cache_stats():
    """{uploads, hashes, tracked_clean, tracked_dirty, tracking_available} of the matrix cache."""
    out = (C.c_longlong * 5)()
    lib().cgb200_cache_stats(out)
    return dict(zip(("uploads", "hashes", "tracked_clean", "tracked_dirty", "tracking_available"), [int(v) for v in out]))


def last_error():
    return lib().cgb200_last_error().decode()


class CgError(RuntimeError):
    pass


def _check(rc, what):
    if rc != 0:
        raise CgError(f"{what} failed: {last_error()}")


class DeviceSystem:
    """A device-resident row block of a system (cgb200_system*)."""

    def __init__(self, handle, what):
        if not handle:
            raise CgError(f"{what} failed: {last_error()}")
        self.h = C.c_void_p(handle)

    @classmethod
    def lap(cls, nx, ny, nz=1, z0=0, nzl=None, bseed=7):
        nzl = nz if nzl is None else nzl
        return cls(lib().cgb200_system_lap(nx, ny, nz, z0, nzl, bseed), "cgb200_system_lap")

    @classmethod
    def dense(cls, n, row0=0, nrows=None, aseed=1, bseed=7):
        nrows = n if nrows is None else nrows
        return cls(lib().cgb200_system_dense(n, row0, nrows, aseed, bseed), "cgb200_system_dense")

    @classmethod
    def from_csr(cls, n_global, row0, row_ptr, cols, vals, b):
        import numpy as np

        row_ptr = np.ascontiguousarray(row_ptr, dtype=np.int32)
        cols = np.ascontiguousarray(cols, dtype=np.int32)
        vals = np.ascontiguousarray(vals, dtype=np.float64)
        b = np.ascontiguousarray(b, dtype=np.float64)
        return cls(
            lib().cgb200_system_from_csr(
                n_global, row0, len(row_ptr) - 1,
                row_ptr.ctypes.data_as(c_int_p), cols.ctypes.data_as(c_int_p),
                vals.ctypes.data_as(c_double_p), b.ctypes.data_as(c_double_p),
            ),
            "cgb200_system_from_csr",
        )

    @property
    def rows(self):
        return lib().cgb200_system_rows(self.h)

    @property
    def nnz(self):
        return lib().cgb200_system_nnz(self.h)

    def solve(self, max_iter, tol=0.0, variant=CLASSIC, check_every=0, jacobi=False, x0=None):
        """x0: None = start from zero (the reference, cg.c:104); True = start from the x the system
        holds (a previous solve's); an array = upload it first.  jacobi: diagonal preconditioner."""
        flags = OPT_JACOBI if jacobi else 0
        if x0 is not None:
            flags |= OPT_X0
            if x0 is not True:
                self.set_x(x0)
        o = Opts(max_iter, tol, variant, check_every, flags)
        r = Result()
        _check(lib().cgb200_system_solve(self.h, C.byref(o), C.byref(r)), "cgb200_system_solve")
        return r

    def solve_async(self, max_iter, variant=CLASSIC, check_every=0):
        """Queue a fixed-count solve; returns a ticket for solve_wait()."""
        o = Opts(max_iter, 0.0, variant, check_every, 0)
        t = C.c_longlong(-1)
        _check(lib().cgb200_system_solve_async(self.h, C.byref(o), C.byref(t)), "cgb200_system_solve_async")
        return t.value

    @staticmethod
    def solve_wait(ticket):
        r = Result()
        _check(lib().cgb200_solve_wait(ticket, C.byref(r)), "cgb200_solve_wait")
        return r

    def gen_rhs(self, seed):
        _check(lib().cgb200_system_gen_rhs(self.h, seed), "cgb200_system_gen_rhs")

    def set_rhs(self, b):
        import numpy as np

        b = np.ascontiguousarray(b, dtype=np.float64)
        assert len(b) == self.rows
        _check(lib().cgb200_system_set_rhs(self.h, b.ctypes.data_as(c_double_p)), "cgb200_system_set_rhs")

    def set_x(self, x0):
        import numpy as np

        x0 = np.ascontiguousarray(x0, dtype=np.float64)
        assert len(x0) == self.rows
        _check(lib().cgb200_system_set_x(self.h, x0.ctypes.data_as(c_double_p)), "cgb200_system_set_x")

    def x(self):
        import numpy as np

        out = np.empty(self.rows, dtype=np.float64)
        _check(lib().cgb200_system_get_x(self.h, out.ctypes.data_as(c_double_p)), "cgb200_system_get_x")
        return out

    def matvec(self, v):
        import numpy as np

        v = np.ascontiguousarray(v, dtype=np.float64)
        assert len(v) == self.rows
        out = np.empty(self.rows, dtype=np.float64)
        _check(
            lib().cgb200_system_matvec(self.h, v.ctypes.data_as(c_double_p), out.ctypes.data_as(c_double_p)),
            "cgb200_system_matvec",
        )
        return out

    def true_relres(self):
        v = C.c_double(0.0)
        _check(lib().cgb200_system_true_relres(self.h, C.byref(v)), "cgb200_system_true_relres")
        return v.value

    def time_kernel(self, which, reps):
        v = C.c_double(0.0)
        _check(lib().cgb200_system_time_kernel(self.h, which, reps, C.byref(v)), "cgb200_system_time_kernel")
        return v.value

    def use_pattern_format(self, on=True):
        """Switch the matvec to the row-pattern dictionary format; True if it applies to this matrix."""
        rc = lib().cgb200_system_use_pattern_format(self.h, 1 if on else 0)
        if rc < 0:
            raise CgError("cgb200_system_use_pattern_format failed: " + last_error())
        return rc == 1

    def use_banded_format(self, on=True):
        """Switch the matvec to diagonal storage; True if the matrix is a mostly full band."""
        rc = lib().cgb200_system_use_banded_format(self.h, 1 if on else 0)
        if rc < 0:
            raise CgError("cgb200_system_use_banded_format failed: " + last_error())
        return rc == 1

    def one_kernel_iteration(self):
        """True when a classic iteration of this system is one cooperative launch (dense path)."""
        return bool(lib().cgb200_system_one_kernel_iteration(self.h))

    def comm_bytes(self):
        """Bytes this rank sends to its peers per classic iteration (0 on one GPU)."""
        return lib().cgb200_system_comm_bytes(self.h)

    def bytes(self):
        a, b = C.c_double(0.0), C.c_double(0.0)
        _check(lib().cgb200_system_bytes(self.h, C.byref(a), C.byref(b)), "cgb200_system_bytes")
        return a.value, b.value

    def free(self):
        if self.h:
            lib().cgb200_system_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass
