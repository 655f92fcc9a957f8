"""Matvec kernel time on long-row CSR systems for the long-row kernels (CGB200_SPMV_LONGROW: the
shuffle-chain k_spmv_rowgroup, the shared-memory k_spmv_rowblock with 4 or 2 warps per CTA) and
for the stream kernel (CGB200_SPMV_ROWGROUP_MIN=0).  One process per
setting (the setting is read once).  GPU box:   python tests/rowgroup_sweep.py > gpurun_out/rowgroup_sweep.log
Uses tests/helpers.py (oracle generators) to build the systems: test infrastructure."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

SYSTEMS = {"course n=52269 345/row": (52269, 172), "banded n=400000 65/row": (400000, 32),
           "banded n=1000000 33/row": (1000000, 16)}

if len(sys.argv) > 1 and sys.argv[1] == "one":
    import __graft_entry__ as g
    import helpers as H

    pkg = g.load_package()
    assert pkg.lib().cgb200_init(-1) == 0
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6550.0
    out = {}
    only = sys.argv[2] if len(sys.argv) > 2 else ""   # e.g. "course": one system (ncu captures)
    for name, (n, hb) in SYSTEMS.items():
        if only and not name.startswith(only):
            continue
        s = H.banded(n, hb)
        d = pkg.DeviceSystem.from_csr(s.n, 0, s.row_ptr, s.cols, s.vals, s.b)
        _, k1 = d.bytes()
        ms = d.time_kernel(1, 50)
        r = d.solve(49)
        out[name] = {"k1_ms": round(ms, 5), "k1_gbs": round(k1 / ms / 1e6, 1), "frac": round(k1 / ms / 1e6 / peak, 3),
                     "iter_ms": round(r.seconds / r.iterations * 1e3, 5)}
        d.free()
    print(json.dumps(out))
else:
    only = sys.argv[1] if len(sys.argv) > 1 else ""
    cfgs = [("group 2,8,256", {"CGB200_SPMV_LONGROW": "group", "CGB200_SPMV_ROWGROUP_MIN": "16"})]
    cfgs += [(f, {"CGB200_SPMV_LONGROW": f, "CGB200_SPMV_ROWGROUP_MIN": "16"})
             for f in ("block", "block16", "block8", "block16x256", "block8x256")]
    cfgs += [("stream", {"CGB200_SPMV_ROWGROUP_MIN": "0"})]
    for label, env in cfgs:
        r = subprocess.run([sys.executable, os.path.abspath(__file__), "one", only], env=dict(os.environ, **env),
                           stdout=subprocess.PIPE, text=True)
        print(label, r.stdout.strip().splitlines()[-1] if r.stdout.strip() else "failed", flush=True)
