"""CPU: the library's reader/writer of the reference's on-disk format (csrc/sysio.cpp)
against the compiled reference's own reader (read_input_file, cg.c:146-218) and numpy.  No GPU."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest

import helpers as H


def _read(pkg, path):
    L = pkg.lib()
    A, b = L.new_mv_struct(), L.new_mv_struct()
    rc = L.cgb200_read_system(str(path).encode(), A, b)
    if rc != 0:
        return rc, pkg.last_error()
    a, bb = A.contents, b.contents
    out = dict(
        n=a.size, nnz=a.nnz, bn=bb.size, bnnz=bb.nnz,
        vals=np.ctypeslib.as_array(a.values, shape=(a.nnz,)).copy() if a.nnz else np.empty(0),
        cols=np.ctypeslib.as_array(a.col_indices, shape=(a.nnz,)).copy() if a.nnz else np.empty(0, np.int32),
        rp=np.ctypeslib.as_array(a.row_ptr, shape=(a.size + 1,)).copy(),
        b=np.ctypeslib.as_array(bb.values, shape=(bb.size,)).copy() if bb.size else np.empty(0),
        b_is_vector=not bb.col_indices and not bb.row_ptr,
    )
    L.cgb200_free_all(A)
    L.cgb200_free_all(b)
    return 0, out


def _write(pkg, s, path, binary):
    A = pkg.MvSparse(s.n, s.nnz, H.dptr(s.vals), H.iptr(s.cols), H.iptr(s.row_ptr))
    b = pkg.MvSparse(s.n, s.n, H.dptr(s.b), None, None)
    assert pkg.lib().cgb200_write_system(str(path).encode(), C.byref(A), C.byref(b), int(binary)) == 0, pkg.last_error()


def _same(got, s):
    assert got["n"] == s.n and got["nnz"] == s.nnz and got["bn"] == s.n and got["bnnz"] == s.n and got["b_is_vector"]
    assert np.array_equal(got["vals"].view(np.uint64), s.vals.view(np.uint64))
    assert np.array_equal(got["cols"], s.cols) and np.array_equal(got["rp"], s.row_ptr)
    assert np.array_equal(got["b"].view(np.uint64), s.b.view(np.uint64))


SYSTEMS = [lambda: H.tridiag4(), lambda: H.lap(32, 32), lambda: H.lap(9, 7, 5), lambda: H.dense(64),
           lambda: H.banded(600, 5), lambda: H.lap(300, 300)]


@pytest.mark.parametrize("mk", SYSTEMS)
@pytest.mark.parametrize("binary", [False, True])
def test_write_then_read_is_bit_exact(pkg, tmp_path, mk, binary):
    s = mk()
    p = tmp_path / ("sys.bin" if binary else "sys.txt")
    _write(pkg, s, p, binary)
    rc, got = _read(pkg, p)
    assert rc == 0, got
    _same(got, s)


@pytest.mark.parametrize("threads", ["1", "3", "16"])
def test_reads_what_the_reference_format_writer_wrote(pkg, tmp_path, threads, monkeypatch):
    # %.17g text as tests/golden/make_golden.py writes it; the piece boundaries move with the thread count
    monkeypatch.setenv("CGB200_IO_THREADS", threads)
    s = H.lap(120, 77)
    p = tmp_path / "in.txt"
    H.write_text_format(s, str(p))
    rc, got = _read(pkg, p)
    assert rc == 0, got
    _same(got, s)


def test_golden_config1_file(pkg):
    s = H.lap(32, 32)
    rc, got = _read(pkg, os.path.join(H.GOLDEN, "config1_lap2d_32x32.txt"))
    assert rc == 0, got
    _same(got, s)


@pytest.mark.skipif(not H.have_ref(), reason="oracle/_ref not built")
def test_matches_the_reference_reader_on_our_own_output(pkg, tmp_path):
    # our writer -> the reference's reader (fresh process: its reader is one-shot, cg.c:235-297) and
    # the reference-format file -> both readers: identical arrays
    s = H.banded(700, 9)
    s.vals[:7] = [1e-310, -0.0, 1.7976931348623157e308, 5e-324, 0.1, 1 / 3, -2.5e-7]   # awkward decimals
    p = tmp_path / "ours.txt"
    _write(pkg, s, p, False)
    code = (
        "import sys, numpy as np; sys.path.insert(0, %r); import helpers as H, ctypes as C\n"
        "R = H.ref(); A = R.new_mv_struct(); b = R.new_mv_struct()\n"
        "assert R.read_input_file(%r.encode(), A, b) == 0\n"
        "a = A.contents\n"
        "np.savez(%r, vals=np.ctypeslib.as_array(a.values, shape=(a.nnz,)), cols=np.ctypeslib.as_array(a.col_indices, shape=(a.nnz,)),\n"
        "         rp=np.ctypeslib.as_array(a.row_ptr, shape=(a.size + 1,)), b=np.ctypeslib.as_array(b.contents.values, shape=(b.contents.size,)))\n"
    ) % (os.path.join(H.ROOT, "tests"), str(p), str(tmp_path / "ref.npz"))
    subprocess.run([sys.executable, "-c", code], check=True)
    ref = np.load(tmp_path / "ref.npz")
    rc, got = _read(pkg, p)
    assert rc == 0, got
    _same(got, s)
    assert np.array_equal(ref["vals"].view(np.uint64), got["vals"].view(np.uint64))
    assert np.array_equal(ref["cols"], got["cols"]) and np.array_equal(ref["rp"], got["rp"])
    assert np.array_equal(ref["b"].view(np.uint64), got["b"].view(np.uint64))


def test_tolerated_spellings(pkg, tmp_path):
    # no final newline (the reference never terminates on it), CR LF, blanks, a trailing comma, '+', exponents,
    # integer-valued reals as indices, text after the fourth line
    p = tmp_path / "t.txt"
    p.write_bytes(b"0, 1,0 ,1.0\r\n0,2, 4\r\n+2,-1e0,-1.0E+0,2.,\r\n 1 ,\t0.5e1")
    rc, got = _read(pkg, p)
    assert rc == 0, got
    assert got["n"] == 2 and got["nnz"] == 4
    assert got["cols"].tolist() == [0, 1, 0, 1] and got["rp"].tolist() == [0, 2, 4]
    assert got["vals"].tolist() == [2.0, -1.0, -1.0, 2.0] and got["b"].tolist() == [1.0, 5.0]
    p.write_text("0\n0,1\n3.5\n7\nthis line and\nthe rest are ignored\n")
    rc, got = _read(pkg, p)
    assert rc == 0 and got["n"] == 1 and got["vals"].tolist() == [3.5] and got["b"].tolist() == [7.0]


@pytest.mark.parametrize("text,why", [
    ("0,1\n0,1\n1.0\n1.0\n", "malformed"),            # 2 column indices, 1 value
    ("0\n0,1\n1.0\n", "4 expected"),                     # a line is missing
    ("0,,1\n0,2\n1,2\n1\n", "token"),                   # empty token between commas
    ("0,x\n0,2\n1,2\n1\n", "token"),                    # not a number
    ("0,1\n0,2\n1,2z\n1\n", "token"),
    ("0,1\n1,3\n1,2\n1\n", "malformed"),                # row pointers do not start at 0
    ("0,1\n0,2\n1,2\n1,2\n", "malformed"),              # rhs longer than n
    ("0,99999999999\n0,2\n1,2\n1\n", "token"),          # index beyond int
    ("", "empty"),
])
def test_errors_are_reported(pkg, tmp_path, text, why):
    p = tmp_path / "bad.txt"
    p.write_text(text)
    rc, msg = _read(pkg, p)
    assert rc == -1 and why in msg, msg


def test_missing_file_and_truncated_binary(pkg, tmp_path):
    rc, msg = _read(pkg, tmp_path / "nope.txt")
    assert rc == -1 and "Failed to open input file" in msg
    s = H.lap(20, 20)
    p = tmp_path / "s.bin"
    _write(pkg, s, p, True)
    raw = p.read_bytes()
    p.write_bytes(raw[: len(raw) - 9])
    rc, msg = _read(pkg, p)
    assert rc == -1 and "truncated" in msg


def test_writer_refuses_non_systems(pkg, tmp_path):
    L = pkg.lib()
    v = np.ones(4)
    vec = pkg.MvSparse(4, 4, H.dptr(v), None, None)
    assert L.cgb200_write_system(str(tmp_path / "x.txt").encode(), C.byref(vec), C.byref(vec), 0) == -1
    s = H.lap(4, 4)
    A = pkg.MvSparse(s.n, s.nnz, H.dptr(s.vals), H.iptr(s.cols), H.iptr(s.row_ptr))
    short = pkg.MvSparse(3, 3, H.dptr(v), None, None)
    assert L.cgb200_write_system(str(tmp_path / "x.txt").encode(), C.byref(A), C.byref(short), 0) == -1


@pytest.mark.parametrize("seed", range(4))
def test_real_spellings_fuzz_against_python_float(pkg, tmp_path, seed):
    # every spelling printf / numpy / MATLAB-style exporters produce for a double must come back as
    # the double Python's float() (correctly rounded, like the reference's strtod) reads from it
    rng = np.random.default_rng(100 + seed)
    n = 1500
    mags = rng.choice([1e-310, 1e-300, 1e-20, 1e-3, 1.0, 1e3, 1e17, 1e300], size=n)
    x = rng.standard_normal(n) * mags
    x[:6] = [0.0, -0.0, 5e-324, 1.7976931348623157e308, 0.1, -1 / 3]
    fmts = ["%.17g", "%.16e", "%.6f", "%r", "%.3g", "%+.17g", "%.17E", "%.0f"]
    toks = []
    for k, v in enumerate(x):
        f = fmts[int(rng.integers(len(fmts)))]
        t = repr(float(v)) if f == "%r" else f % v
        if len(t) > 400:
            t = "%.17g" % v
        pad_l, pad_r = " " * int(rng.integers(0, 3)), " " * int(rng.integers(0, 3))
        toks.append(pad_l + t + pad_r)
    want = np.array([float(t) for t in toks])
    # a 1-row "matrix" whose values line carries the tokens; b holds them once more on the fourth line
    cols = ",".join(str(i) for i in range(n))
    p = tmp_path / "fuzz.txt"
    p.write_text(f"{cols}\n0,{n}\n{','.join(toks)}\n{toks[0]}\n")
    for threads in ("1", "7"):
        os.environ["CGB200_IO_THREADS"] = threads
        try:
            rc, got = _read(pkg, p)
        finally:
            del os.environ["CGB200_IO_THREADS"]
        assert rc == 0, got
        assert got["n"] == 1 and got["nnz"] == n
        assert np.array_equal(got["vals"].view(np.uint64), want.view(np.uint64))
        assert got["b"].view(np.uint64)[0] == want.view(np.uint64)[0]


@pytest.mark.skipif(not H.have_ref(), reason="oracle/_ref not built")
def test_reference_cli_solves_a_file_our_writer_wrote(pkg, tmp_path):
    # golden input -> our reader -> our writer -> the unmodified reference command line: the stdout the
    # reference printed for the golden input itself (tests/golden/config1_cg_stdout.txt)
    L = pkg.lib()
    A, b = L.new_mv_struct(), L.new_mv_struct()
    assert L.cgb200_read_system(os.path.join(H.GOLDEN, "config1_lap2d_32x32.txt").encode(), A, b) == 0
    p = tmp_path / "rewritten.txt"
    assert L.cgb200_write_system(str(p).encode(), A, b, 0) == 0, pkg.last_error()
    L.cgb200_free_all(A)
    L.cgb200_free_all(b)
    out = subprocess.run([H.REF_CG, str(p), "30"], stdout=subprocess.PIPE, text=True, check=True).stdout

    def x_lines(txt):   # the printed vector (mv_ops.c:92-93); not the timing line or the pointer
        return [l for l in txt.splitlines() if l.startswith("\t") and ":" not in l]

    want = x_lines(open(os.path.join(H.GOLDEN, "config1_cg_stdout.txt")).read())
    assert len(want) == 1024 and x_lines(out) == want
