"""Per-kernel device times and roofline fractions for the BASELINE workloads (1 GPU)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as g

pkg = g.load_package()
L = pkg.lib()
assert L.cgb200_init(-1) == 0, pkg.last_error()
peak = 6551.7
try:
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass

which = sys.argv[1:] or ["lap2d", "lap3d384", "lap3d512", "dense"]
for w in which:
    if w == "lap2d":
        d = pkg.DeviceSystem.lap(4096, 4096, 1)
    elif w == "lap3d384":
        d = pkg.DeviceSystem.lap(384, 384, 384)
    elif w == "lap3d512":
        d = pkg.DeviceSystem.lap(512, 512, 512)
    elif w == "dense":
        d = pkg.DeviceSystem.dense(32768)
    elif w == "dense8k":
        d = pkg.DeviceSystem.dense(8192)
    else:
        raise SystemExit("unknown " + w)
    n = d.rows
    it_b, k1_b = d.bytes()
    t1 = d.time_kernel(1, 20)
    t2 = d.time_kernel(2, 20)
    t3 = d.time_kernel(3, 20)
    r = d.solve(199)
    ms_it = r.seconds * 1e3 / r.iterations
    print(json.dumps({
        "workload": w, "n": n, "nnz": d.nnz,
        "K1_ms": round(t1, 4), "K1_GBs": round(k1_b / t1 / 1e6, 1), "K1_frac": round(k1_b / t1 / 1e6 / peak, 3),
        "K2_ms": round(t2, 4), "K2_GBs": round(48 * n / t2 / 1e6, 1),
        "K3_ms": round(t3, 4), "K3_GBs": round(24 * n / t3 / 1e6, 1),
        "iters": r.iterations, "status": r.status, "ms_per_iter": round(ms_it, 4),
        "it_per_s": round(1e3 / ms_it, 1), "iter_GBs": round(it_b / ms_it / 1e6, 1),
        "iter_frac": round(it_b / ms_it / 1e6 / peak, 3), "relres": r.relres,
    }), flush=True)
    d.free()
