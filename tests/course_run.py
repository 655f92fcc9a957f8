"""The reference's own "full" run (Makefile:27: `./cg input/Ab.txt 30`) on a stand-in for the
course matrix that is not shipped: banded SPD, n = 52269, 345 entries per row,
18.0 M non-zeros (cg.c:231-232,260), written in the reference's text format.

    python tests/course_run.py ref  > profiles/<round>_course_reference.json   (CPU, ~11 min: here)
    python tests/course_run.py b200 > profiles/<round>_course_b200.json        (GPU box)

`ref`  : the unmodified reference from oracle/_ref -- its reader on the file, then its conj_grad
         (30 iterations = 31 O(n^2) matvecs), timed; x compared BITWISE with the oracle restatement
         (which extends the oracle's pinning to a system of the reference's own scale).
`b200` : the same command line on the product (conjugate-gradient_b200/cg <file> 30 y), wall
         clock from exec to exit -- process start, CUDA context, file parse, upload, 31 iterations,
         download -- plus the same steps timed in-process; x compared with the oracle (<= 1e-8).
Test infrastructure: the oracle and the compiled reference are used as checker and timed baseline.
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

N, HB, M = 52269, 172, 30


def main():
    mode = sys.argv[1] if len(sys.argv) > 1 else "b200"
    n = int(sys.argv[2]) if len(sys.argv) > 2 else N
    hb = int(sys.argv[3]) if len(sys.argv) > 3 else HB
    pkg = g.load_package()
    L = pkg.lib()
    s = H.banded(n, hb)
    out = {"system": f"banded SPD n={s.n}, half bandwidth {hb}, nnz={s.nnz}", "max_iter": M,
           "command": f"cg <file> {M}", "host_threads": os.cpu_count()}
    t = time.perf_counter()
    xo, it, _, _ = H.oracle_cg(s, M)
    out["oracle_seconds"] = round(time.perf_counter() - t, 3)
    with tempfile.TemporaryDirectory(dir=os.environ.get("TMPDIR", "/tmp")) as d:
        path = os.path.join(d, "Ab.txt")
        A = pkg.MvSparse(s.n, s.nnz, H.dptr(s.vals), H.iptr(s.cols), H.iptr(s.row_ptr))
        b = pkg.MvSparse(s.n, s.n, H.dptr(s.b), None, None)
        assert L.cgb200_write_system(path.encode(), C.byref(A), C.byref(b), 0) == 0, pkg.last_error()
        out["file_mb"] = round(os.path.getsize(path) / 1e6, 1)
        if mode == "ref":
            R = H.ref()
            RA, Rb = R.new_mv_struct(), R.new_mv_struct()
            t = time.perf_counter()
            assert R.read_input_file(path.encode(), RA, Rb) == 0
            t_read = time.perf_counter() - t
            xp = H.MvSparseP()
            t = time.perf_counter()
            assert R.conj_grad(M, RA, Rb, C.byref(xp)) == 0
            t_solve = time.perf_counter() - t
            x = np.ctypeslib.as_array(xp.contents.values, shape=(s.n,)).copy()
            out.update({"impl": "reference (oracle/_ref, gcc -g as its Makefile), 1 core",
                        "read_seconds": round(t_read, 3), "solve_seconds": round(t_solve, 3),
                        "total_seconds": round(t_read + t_solve, 3), "iterations": M + 1,
                        "it_per_s": round((M + 1) / t_solve, 4),
                        "x_bitwise_equal_oracle": bool(np.array_equal(x.view(np.uint64), xo.view(np.uint64)))})
        else:
            assert L.cgb200_device_available() == 1, "no CUDA device"
            # the command a user of the reference types
            runs = []
            for _ in range(3):
                t = time.perf_counter()
                r = subprocess.run([pkg.CG_DRIVER, path, str(M), "y"], stdout=subprocess.PIPE, text=True)
                runs.append(time.perf_counter() - t)
                assert r.returncode == 0, r.stdout[-500:]
            out["cli_wall_seconds"] = [round(v, 3) for v in runs]
            # the same steps in-process
            assert L.cgb200_init(-1) == 0, pkg.last_error()
            A2, b2 = L.new_mv_struct(), L.new_mv_struct()
            t = time.perf_counter()
            assert L.cgb200_read_system(path.encode(), A2, b2) == 0, pkg.last_error()
            t_read = time.perf_counter() - t
            xp = pkg.MvSparseP()
            t = time.perf_counter()
            assert L.conj_grad(M, A2, b2, C.byref(xp)) == 0, pkg.last_error()
            t_first = time.perf_counter() - t      # uploads A
            x = np.ctypeslib.as_array(xp.contents.values, shape=(s.n,)).copy()
            L.cgb200_free_all(xp)
            xp = pkg.MvSparseP()
            res = pkg.Result()
            o = pkg.Opts(M, 0.0, 0, 0, 0)
            t = time.perf_counter()
            assert L.cgb200_solve(A2, b2, C.byref(xp), C.byref(o), C.byref(res)) == 0, pkg.last_error()
            t_warm = time.perf_counter() - t       # A already resident
            rel = float(np.linalg.norm(x - xo) / np.linalg.norm(xo))
            out.update({"impl": "libcgb200 on one B200", "read_seconds": round(t_read, 3),
                        "conj_grad_seconds_incl_upload": round(t_first, 4),
                        "conj_grad_seconds_matrix_resident": round(t_warm, 5),
                        "device_loop_seconds": round(res.seconds, 6), "iterations": int(res.iterations),
                        "it_per_s_device": round(res.iterations / res.seconds, 1),
                        "x_rel_diff_vs_oracle": rel, "parity_ok": bool(rel <= 1e-8 and res.iterations == it)})
    print(json.dumps(out))


if __name__ == "__main__":
    main()
