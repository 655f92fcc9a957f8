"""Throughput of the library's system reader/writer (csrc/sysio.cpp) beside the
reference's own reader (read_input_file, cg.c:146-218, from oracle/_ref/libcgref.so), on the
course-sized system the reference's reader is pre-sized for (cg.c:231-232,260: about 18.02 M
non-zeros, n = 52269): a banded SPD matrix, n = 52269, 345 entries per row.  Host only -- runs
wherever oracle/_ref exists (this container); no GPU.

    python tests/io_bench.py [n half_bandwidth] > profiles/<round>_io_bench.json

Checker use of oracle/: the reference reader is the baseline being timed, and its arrays are
compared bit for bit with the new reader's.
"""
import ctypes as C
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as g  # noqa: E402
import helpers as H  # noqa: E402

REF_CODE = r"""
import sys, time, numpy as np
sys.path.insert(0, %(tests)r)
import helpers as H, ctypes as C
R = H.ref(); A = R.new_mv_struct(); b = R.new_mv_struct()
t = time.perf_counter()
rc = R.read_input_file(%(path)r.encode(), A, b)
dt = time.perf_counter() - t
assert rc == 0
a = A.contents
np.savez(%(out)r, seconds=dt, vals=np.ctypeslib.as_array(a.values, shape=(a.nnz,)),
         cols=np.ctypeslib.as_array(a.col_indices, shape=(a.nnz,)),
         rp=np.ctypeslib.as_array(a.row_ptr, shape=(a.size + 1,)),
         b=np.ctypeslib.as_array(b.contents.values, shape=(b.contents.size,)))
"""


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 52269
    hb = int(sys.argv[2]) if len(sys.argv) > 2 else 172
    pkg = g.load_package()
    L = pkg.lib()
    s = H.banded(n, hb)
    A = pkg.MvSparse(s.n, s.nnz, H.dptr(s.vals), H.iptr(s.cols), H.iptr(s.row_ptr))
    b = pkg.MvSparse(s.n, s.n, H.dptr(s.b), None, None)
    out = {"system": f"banded SPD n={s.n}, half bandwidth {hb}", "n": s.n, "nnz": s.nnz,
           "host_threads": os.cpu_count()}
    with tempfile.TemporaryDirectory(dir=os.environ.get("TMPDIR", "/tmp")) as d:
        txt, binf = os.path.join(d, "sys.txt"), os.path.join(d, "sys.bin")

        def timed(fn, reps=3):
            best = 1e99
            for _ in range(reps):
                t = time.perf_counter()
                fn()
                best = min(best, time.perf_counter() - t)
            return best

        def read(path):
            a_, b_ = L.new_mv_struct(), L.new_mv_struct()
            assert L.cgb200_read_system(path.encode(), a_, b_) == 0, pkg.last_error()
            return a_, b_

        def read_and_free(path):
            a_, b_ = read(path)
            L.cgb200_free_all(a_)
            L.cgb200_free_all(b_)

        tw = timed(lambda: L.cgb200_write_system(txt.encode(), C.byref(A), C.byref(b), 0), 2)
        mb = os.path.getsize(txt) / 1e6
        out["text_file_mb"] = round(mb, 1)
        out["write_text"] = {"seconds": round(tw, 3), "mb_per_s": round(mb / tw, 1)}
        tr = timed(lambda: read_and_free(txt))
        out["read_text"] = {"seconds": round(tr, 3), "mb_per_s": round(mb / tr, 1), "threads": "all"}
        os.environ["CGB200_IO_THREADS"] = "1"
        tr1 = timed(lambda: read_and_free(txt), 2)
        del os.environ["CGB200_IO_THREADS"]
        out["read_text_1_thread"] = {"seconds": round(tr1, 3), "mb_per_s": round(mb / tr1, 1)}

        twb = timed(lambda: L.cgb200_write_system(binf.encode(), C.byref(A), C.byref(b), 1), 2)
        mbb = os.path.getsize(binf) / 1e6
        trb = timed(lambda: read_and_free(binf))
        out["binary_file_mb"] = round(mbb, 1)
        out["write_binary"] = {"seconds": round(twb, 3), "mb_per_s": round(mbb / twb, 1)}
        out["read_binary"] = {"seconds": round(trb, 3), "mb_per_s": round(mbb / trb, 1)}

        # bit-exactness of both paths against the arrays we started from
        for path in (txt, binf):
            a_, b_ = read(path)
            ac = a_.contents
            assert ac.size == s.n and ac.nnz == s.nnz
            assert np.array_equal(np.ctypeslib.as_array(ac.values, shape=(s.nnz,)).view(np.uint64), s.vals.view(np.uint64))
            assert np.array_equal(np.ctypeslib.as_array(ac.col_indices, shape=(s.nnz,)), s.cols)
            assert np.array_equal(np.ctypeslib.as_array(ac.row_ptr, shape=(s.n + 1,)), s.row_ptr)
            assert np.array_equal(np.ctypeslib.as_array(b_.contents.values, shape=(s.n,)).view(np.uint64), s.b.view(np.uint64))
            L.cgb200_free_all(a_)
            L.cgb200_free_all(b_)
        out["round_trip_bit_exact"] = True

        if H.have_ref():
            npz = os.path.join(d, "ref.npz")
            code = REF_CODE % {"tests": os.path.join(ROOT, "tests"), "path": txt, "out": npz}
            subprocess.run([sys.executable, "-c", code], check=True)
            r = np.load(npz)
            same = (np.array_equal(r["vals"].view(np.uint64), s.vals.view(np.uint64)) and np.array_equal(r["cols"], s.cols)
                    and np.array_equal(r["rp"], s.row_ptr) and np.array_equal(r["b"].view(np.uint64), s.b.view(np.uint64)))
            rs = float(r["seconds"])
            out["reference_reader"] = {"seconds": round(rs, 3), "mb_per_s": round(mb / rs, 1), "threads": 1,
                                       "what": "read_input_file of the unmodified reference (oracle/_ref, gcc -g as its Makefile)",
                                       "same_arrays_as_ours": bool(same)}
            out["speedup_vs_reference_reader"] = round(rs / tr, 1)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
