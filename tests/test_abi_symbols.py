"""CPU: libcgb200.so loads without a GPU and exports every symbol include/*.h declare; the
host-only parts of the reference surface behave like the reference; arithmetic entry points fail
loudly (no CPU fallback) when there is no device."""
import ctypes as C
import io
import os
import re
import subprocess

import numpy as np
import pytest

import helpers as H

ROOT = H.ROOT


def declared_functions(header):
    txt = open(os.path.join(ROOT, "include", header)).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    names = re.findall(r"\b([A-Za-z_][A-Za-z0-9_]*)\s*\([^;{]*\)\s*;", txt)
    return sorted(set(n for n in names if n not in ("defined",)))


def test_headers_declare_the_reference_surface():
    fns = declared_functions("mv_ops.h")
    # reference mv_ops.h:26-42 + cg.c:24
    for n in ["new_mv_struct", "new_mv_struct_with_size", "free_mv_struct", "mv_deep_copy", "print_sparse",
              "mat_get_row", "dot_product", "sv_mult", "mv_mult", "vec_add", "vec_sub", "conj_grad"]:
        assert n in fns


def test_library_exports_every_declared_symbol(pkg):
    L = pkg.lib()
    out = subprocess.run(["nm", "-D", "--defined-only", pkg.LIB_PATH], stdout=subprocess.PIPE, text=True, check=True).stdout
    exported = {ln.split()[-1] for ln in out.splitlines() if " T " in ln}
    for hdr in ("mv_ops.h", "cgb200.h"):
        for fn in declared_functions(hdr):
            assert fn in exported, f"{fn} declared in include/{hdr} but not exported"
            getattr(L, fn)


def test_struct_layout_is_the_reference_abi(pkg):
    # mv_ops.h:17-23 on LP64: int, int, double*, int*, int* = 32 bytes
    assert C.sizeof(pkg.MvSparse) == 32
    assert pkg.MvSparse.values.offset == 8 and pkg.MvSparse.col_indices.offset == 16 and pkg.MvSparse.row_ptr.offset == 24


def test_lifecycle_is_host_side_and_matches_reference_behaviour(pkg):
    L = pkg.lib()
    v = L.new_mv_struct_with_size(5)
    assert v.contents.size == 5 and v.contents.nnz == 5
    assert not v.contents.col_indices and not v.contents.row_ptr
    assert np.array_equal(np.ctypeslib.as_array(v.contents.values, shape=(5,)), np.zeros(5))
    z = L.new_mv_struct()
    assert z.contents.size == 0 and not z.contents.values
    s = H.lap(4, 3)
    A = pkg.MvSparse(s.n, s.nnz, H.dptr(s.vals), H.iptr(s.cols), H.iptr(s.row_ptr))
    cp = L.mv_deep_copy(C.byref(A))
    assert cp.contents.size == s.n and cp.contents.nnz == s.nnz
    assert np.array_equal(np.ctypeslib.as_array(cp.contents.values, shape=(s.nnz,)), s.vals)
    assert np.array_equal(np.ctypeslib.as_array(cp.contents.col_indices, shape=(s.nnz,)), s.cols)
    assert np.array_equal(np.ctypeslib.as_array(cp.contents.row_ptr, shape=(s.n + 1,)), s.row_ptr)
    assert C.addressof(cp.contents.values.contents) != s.vals.ctypes.data
    assert not L.mv_deep_copy(None)
    L.cgb200_free_all(cp)
    L.free_mv_struct(z)
    L.free_mv_struct(None)


def test_mat_get_row(pkg):
    L = pkg.lib()
    s = H.lap(4, 3)
    A = pkg.MvSparse(s.n, s.nnz, H.dptr(s.vals), H.iptr(s.cols), H.iptr(s.row_ptr))
    row = np.full(s.n, 7.0)
    for i in range(s.n):
        assert L.mat_get_row(C.byref(A), i, H.dptr(row)) == 0
        want = np.zeros(s.n)
        want[s.cols[s.row_ptr[i]:s.row_ptr[i + 1]]] = s.vals[s.row_ptr[i]:s.row_ptr[i + 1]]
        assert np.array_equal(row, want)
    assert L.mat_get_row(None, 0, H.dptr(row)) == -1            # mv_ops.c:103-104
    assert L.mat_get_row(C.byref(A), 0, None) == -1
    # unlike the reference's scan a diagonal matrix does not bleed into the next row
    n = 4
    rp = np.arange(n + 1, dtype=np.int32)
    ci = np.arange(n, dtype=np.int32)
    va = np.array([1.0, 2.0, 3.0, 4.0])
    D = pkg.MvSparse(n, n, H.dptr(va), H.iptr(ci), H.iptr(rp))
    r = np.empty(n)
    assert L.mat_get_row(C.byref(D), 0, H.dptr(r)) == 0
    assert np.array_equal(r, [1, 0, 0, 0])


def test_argument_errors_need_no_device(pkg):
    L = pkg.lib()
    a, b = H.rhs(8, 1), H.rhs(9, 2)
    va = pkg.MvSparse(8, 8, H.dptr(a), None, None)
    vb = pkg.MvSparse(9, 9, H.dptr(b), None, None)
    out = pkg.MvSparseP()
    assert L.vec_add(C.byref(va), C.byref(vb), C.byref(out)) == -1
    assert L.vec_sub(None, C.byref(vb), C.byref(out)) == -1
    assert L.sv_mult(2.0, None, C.byref(out)) == -1
    assert L.dot_product(C.byref(va), C.byref(vb)) == -1.0
    assert L.dot_product(None, None) == -1.0
    assert L.mv_mult(None, C.byref(va), C.byref(out)) == -1
    assert L.conj_grad(3, None, C.byref(va), C.byref(out)) == -1


def test_no_cpu_fallback(pkg):
    """Without a device the arithmetic fails loudly; with one this test has nothing to prove."""
    L = pkg.lib()
    if L.cgb200_device_available():
        pytest.skip("a CUDA device is present")
    a = H.rhs(8, 1)
    va = pkg.MvSparse(8, 8, H.dptr(a), None, None)
    out = pkg.MvSparseP()
    assert L.sv_mult(2.0, C.byref(va), C.byref(out)) == -1
    assert "no CPU fallback" in pkg.last_error()
    assert L.dot_product(C.byref(va), C.byref(va)) == -1.0
    s = H.tridiag4()
    A = pkg.MvSparse(s.n, s.nnz, H.dptr(s.vals), H.iptr(s.cols), H.iptr(s.row_ptr))
    b = pkg.MvSparse(s.n, s.n, H.dptr(s.b), None, None)
    assert L.conj_grad(1, C.byref(A), C.byref(b), C.byref(out)) == -1
    with pytest.raises(pkg.CgError):
        pkg.DeviceSystem.lap(8, 8)


def test_product_sources_never_touch_the_oracle():
    """The product path may not import, link or execute anything under oracle/ (tier rule 3)."""
    pk = os.path.join(ROOT, "conjugate-gradient_b200")
    bad = []
    for dp, _, fns in os.walk(pk):
        for fn in fns:
            if fn.endswith((".cu", ".cuh", ".h", ".c", ".cpp", ".py", "Makefile")):
                txt = open(os.path.join(dp, fn), errors="ignore").read()
                if re.search(r"oracle/|liboracle|cg_oracle|libcgref", txt):
                    bad.append(os.path.join(dp, fn))
    assert not bad, bad
    ldd = subprocess.run(["ldd", os.path.join(pk, "libcgb200.so")], stdout=subprocess.PIPE, text=True).stdout
    assert "oracle" not in ldd and "cgref" not in ldd


def test_content_hash_sees_any_single_edit_and_ignores_thread_count(pkg, monkeypatch):
    """The strict matrix-cache policy (include/cgb200.h) rests on this hash: host only."""
    L = pkg.lib()
    s = H.banded(3000, 40)      # ~240k entries: several words per hashing lane, one 4 MiB piece
    A = pkg.MvSparse(s.n, s.nnz, H.dptr(s.vals), H.iptr(s.cols), H.iptr(s.row_ptr))
    h0 = L.cgb200_content_hash(C.byref(A))
    assert h0 != 0 and L.cgb200_content_hash(C.byref(A)) == h0
    rng = np.random.default_rng(5)
    for k in list(rng.integers(0, s.nnz, 40)) + [0, s.nnz - 1]:
        old = s.vals[k]
        s.vals[k] = np.nextafter(old, 2 * old + 1)      # one ulp
        assert L.cgb200_content_hash(C.byref(A)) != h0, k
        s.vals[k] = old
    for k in (0, s.nnz // 3, s.nnz - 1):
        s.cols[k] ^= 1
        assert L.cgb200_content_hash(C.byref(A)) != h0
        s.cols[k] ^= 1
    s.row_ptr[s.n // 2] += 1
    assert L.cgb200_content_hash(C.byref(A)) != h0
    s.row_ptr[s.n // 2] -= 1
    assert L.cgb200_content_hash(C.byref(A)) == h0
    # a matrix larger than one piece: the same hash with 1 and with many threads
    big = H.lap(700, 700)       # 2.4 M entries: 19 MB of values, 5 pieces
    B = pkg.MvSparse(big.n, big.nnz, H.dptr(big.vals), H.iptr(big.cols), H.iptr(big.row_ptr))
    monkeypatch.setenv("CGB200_HASH_THREADS", "1")
    h1 = L.cgb200_content_hash(C.byref(B))
    monkeypatch.setenv("CGB200_HASH_THREADS", "7")
    assert L.cgb200_content_hash(C.byref(B)) == h1
    big.vals[big.nnz - 5] = 3.0
    assert L.cgb200_content_hash(C.byref(B)) != h1
    assert L.cgb200_content_hash(None) == 0


def test_write_tracking_probe_needs_no_gpu(pkg):
    # the default matrix-cache policy asks the kernel whether the caller's arrays were written
    # (userfaultfd async write-protection + PAGEMAP_SCAN); where that is missing it hashes instead
    L = pkg.lib()
    a = L.cgb200_write_tracking_available()
    assert a in (0, 1) and L.cgb200_write_tracking_available() == a


def test_write_tracking_sees_writes_and_only_writes(pkg):
    # csrc/writewatch.cpp on a numpy array, no GPU involved: a read is not a write, one stored byte
    # anywhere in the interior is, clearing the marks starts over, an unwatched range says "cannot tell"
    import numpy as np

    L = pkg.lib()
    if not L.cgb200_write_tracking_available():
        pytest.skip("this kernel has no userfaultfd async write protection / PAGEMAP_SCAN")
    a = np.ones(3 << 20)                         # 24 MiB, touched
    p, nb = a.ctypes.data, a.nbytes
    op = L.cgb200_write_tracking_op
    assert op(p, nb, 1) == -1                    # not watched yet
    assert op(p, nb, 0) == 0
    assert op(p, nb, 1) == 0
    assert float(a[12345:99999].sum()) == 99999 - 12345 and op(p, nb, 1) == 0
    a[(1 << 20) + 17] = 2.0
    assert op(p, nb, 1) == 1
    assert op(p, nb, 2) == 0 and op(p, nb, 1) == 0
    a[-1000] = 3.0                               # the last full page of the interior
    assert op(p, nb, 1) == 1
    assert op(p, nb, 2) == 0
    a[5000:5010] = a[5000:5010].copy()           # the same values are still a write
    assert op(p, nb, 1) == 1
    b = np.zeros(1 << 20)                        # a second, independent range
    assert op(b.ctypes.data, b.nbytes, 0) == 0 and op(b.ctypes.data, b.nbytes, 1) == 0
    assert op(p, nb, 3) == 0 and op(p, nb, 1) == -1
    assert op(b.ctypes.data, b.nbytes, 1) == 0
    b[7700] = 1.0                                # (not in the array's first, partial page: only whole pages are watched)
    assert op(b.ctypes.data, b.nbytes, 1) == 1
    assert op(b.ctypes.data, b.nbytes, 3) == 0
    assert op(p, 100, 0) == -1                   # no whole page inside


def test_write_tracking_on_transparent_huge_pages(pkg):
    # multi-GB matrices are backed by 2 MiB pages where the system allows it: one stored byte inside a
    # huge page is seen, re-arming clears it, and the untouched huge pages stay clean
    import ctypes as C

    import numpy as np

    L = pkg.lib()
    if not L.cgb200_write_tracking_available():
        pytest.skip("this kernel has no userfaultfd async write protection / PAGEMAP_SCAN")
    a = np.empty(4 << 20)                        # 32 MiB, untouched
    lo = (a.ctypes.data + 4095) & ~4095
    ln = (a.ctypes.data + a.nbytes - lo) & ~4095
    C.CDLL(None, use_errno=True).madvise(C.c_void_p(lo), C.c_size_t(ln), 14)    # MADV_HUGEPAGE
    a[:] = 1.0                                   # first touch: huge pages where available
    huge_kb = 0
    with open("/proc/self/smaps") as f:
        inside = False
        for ln_ in f:
            if "-" in ln_.split(" ")[0] and ln_[0] in "0123456789abcdef":
                b0, b1 = (int(x, 16) for x in ln_.split(" ")[0].split("-"))
                inside = b0 <= lo + (8 << 20) < b1       # (the madvise split the mapping: look in its middle)
            elif inside and ln_.startswith("AnonHugePages:"):
                huge_kb = int(ln_.split()[1])
    op = L.cgb200_write_tracking_op
    p, nb = a.ctypes.data, a.nbytes
    assert op(p, nb, 0) == 0 and op(p, nb, 1) == 0
    for k in (300_000, 1_500_000, 2_621_440 + 7, a.size - 600):
        assert float(a[k - 100:k + 100].sum()) == 200.0 and op(p, nb, 1) == 0     # reads leave it clean
        a[k] = 2.0
        assert op(p, nb, 1) == 1, (k, huge_kb)
        assert op(p, nb, 2) == 0 and op(p, nb, 1) == 0, (k, huge_kb)
        a[k] = 1.0
        assert op(p, nb, 1) == 1
        assert op(p, nb, 2) == 0 and op(p, nb, 1) == 0
    assert op(p, nb, 3) == 0
    print("AnonHugePages under the array:", huge_kb, "kB")



def test_every_environment_variable_the_library_reads_is_documented():
    # INTEGRATION.md section 7 is the list a maintainer of the reference looks at
    import glob
    import re

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    pk = os.path.join(root, "conjugate-gradient_b200")
    read = set()
    for f in glob.glob(os.path.join(pk, "csrc", "*")) + glob.glob(os.path.join(pk, "driver", "*")) + [os.path.join(pk, "__init__.py")]:
        with open(f, errors="replace") as fh:
            t = fh.read()
        read |= set(re.findall(r'getenv\("(CGB200_[A-Z0-9_]+)"\)', t))
        read |= set(re.findall(r'environ\.get\("(CGB200_[A-Z0-9_]+)"', t))
    with open(os.path.join(root, "INTEGRATION.md")) as fh:
        doc = fh.read()
    assert len(read) > 20
    assert sorted(e for e in read if e not in doc) == []
