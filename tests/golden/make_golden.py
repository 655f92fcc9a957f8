"""Generates tests/golden/*.npz by running the UNMODIFIED reference (oracle/_ref, built from
/root/reference by oracle/Makefile) in this container.  The reference ships no test vectors of its
own, so these fixtures pin the oracle and the CUDA path to the reference's
actual outputs.  Re-run only where /root/reference is mounted:

    make -C oracle && python tests/golden/make_golden.py
"""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import helpers as H  # noqa: E402

assert H.have_ref(), "oracle/_ref/libcgref.so missing: run `make -C oracle` where /root/reference exists"
R = H.ref()


def ref_ops(sysm, seed):
    """mv_mult / dot_product / sv_mult / vec_add / vec_sub of the reference on seeded vectors."""
    n = sysm.n
    u, v = H.rhs(n, seed), H.rhs(n, seed + 1)
    su = H.MvSparse(n, n, H.dptr(u), None, None)
    sv = H.MvSparse(n, n, H.dptr(v), None, None)
    A = sysm.A_struct()
    out = {}
    for name, call in [
        ("mv_mult", lambda o: R.mv_mult(C.byref(A), C.byref(su), C.byref(o))),
        ("sv_mult", lambda o: R.sv_mult(C.c_double(-1.7), C.byref(su), C.byref(o))),
        ("vec_add", lambda o: R.vec_add(C.byref(su), C.byref(sv), C.byref(o))),
        ("vec_sub", lambda o: R.vec_sub(C.byref(su), C.byref(sv), C.byref(o))),
    ]:
        o = H.MvSparseP()
        assert call(o) == 0
        out[name] = np.ctypeslib.as_array(o.contents.values, shape=(n,)).copy()
    out["dot"] = np.array([R.dot_product(C.byref(su), C.byref(sv))])
    out["u"], out["v"] = u, v
    return out


CASES = [
    ("tridiag4_m0", H.tridiag4, 0),
    ("tridiag4_m1", H.tridiag4, 1),
    ("lap2d_32x32_m30", lambda: H.lap(32, 32), 30),     # config 1 stand-in: ./cg file 30
    ("lap2d_17x9_m12", lambda: H.lap(17, 9), 12),
    ("lap3d_12_m20", lambda: H.lap(12, 12, 12), 20),
    ("lap3d_9x7x5_m10", lambda: H.lap(9, 7, 5), 10),
    ("banded_600_5_m15", lambda: H.banded(600, 5), 15),
    ("banded_300_40_m8", lambda: H.banded(300, 40), 8),
    ("dense_64_m3", lambda: H.dense(64), 3),
    ("dense_257_m4", lambda: H.dense(257), 4),
]

for name, mk, M in CASES:
    s = mk()
    x = H.ref_cg(s, M)
    assert np.all(np.isfinite(x)), name
    ops = ref_ops(s, 40)
    np.savez_compressed(
        os.path.join(HERE, name + ".npz"), row_ptr=s.row_ptr, cols=s.cols, vals=s.vals, b=s.b, max_iter=M, x=x,
        **{"op_" + k: v for k, v in ops.items()},
    )
    print(name, "n", s.n, "nnz", s.nnz, "|x|", float(np.linalg.norm(x)))

# config 1 through the reference's own CLI: input file in its 4-line text format and its stdout
s = H.lap(32, 32)
inp = os.path.join(HERE, "config1_lap2d_32x32.txt")
H.write_text_format(s, inp)
out = subprocess.run([H.REF_CG, inp, "30"], stdout=subprocess.PIPE, text=True, check=True).stdout
lines = [ln for ln in out.splitlines() if not ln.startswith("\tValues:") and not ln.startswith("CG took")]
with open(os.path.join(HERE, "config1_cg_stdout.txt"), "w") as f:
    f.write("\n".join(lines) + "\n")
print("config1 CLI:", len(lines), "lines")
