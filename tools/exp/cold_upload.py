"""Where does the first call's time go?  Dense n = 32768 struct in pageable memory; first mv_mult under
each cache policy, staged upload on / off.  python tools/exp/cold_upload.py   (one GPU)"""
import ctypes as C
import importlib
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
pkg = importlib.import_module("conjugate-gradient_b200")
L = pkg.lib()
n = int(os.environ.get("N", 32768))
nnz = n * n
vals = np.random.default_rng(1).random(nnz)
cols = np.tile(np.arange(n, dtype=np.int32), n)
rp = (np.arange(n + 1, dtype=np.int64) * n).astype(np.int32)
v = np.ones(n)
A = pkg.MvSparse(n, nnz, vals.ctypes.data_as(pkg.c_double_p), cols.ctypes.data_as(pkg.c_int_p), rp.ctypes.data_as(pkg.c_int_p))
vv = pkg.MvSparse(n, n, v.ctypes.data_as(pkg.c_double_p), None, None)
out = pkg.MvSparseP()
L.mv_mult(C.byref(A), C.byref(vv), C.byref(out))      # context, allocator warm
L.cgb200_invalidate(C.byref(A))
for pol, name in ((pkg.CACHE_OFF, "off"), (pkg.CACHE_STRICT, "strict"), (pkg.CACHE_TRACKED, "tracked")):
    for staged in ("1", "0"):
        os.environ["CGB200_STAGED_UPLOAD"] = staged
        L.cgb200_set_matrix_cache(pol)
        t = time.perf_counter()
        assert L.mv_mult(C.byref(A), C.byref(vv), C.byref(out)) == 0
        t1 = time.perf_counter() - t
        t = time.perf_counter()
        assert L.mv_mult(C.byref(A), C.byref(vv), C.byref(out)) == 0
        t2 = time.perf_counter() - t
        print(f"policy {name:8s} staged {staged}: first call {t1*1e3:8.1f} ms   second {t2*1e3:8.1f} ms", flush=True)
        L.cgb200_invalidate(C.byref(A))
