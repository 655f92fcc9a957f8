"""The long-row workload of bench.py alone (banded SPD n = 52269, 345 entries per row), for A/B runs
of the long-row matvec under different environment settings:
    CGB200_L2_RESIDENT_MB=0 python tools/longrow_bench.py
Prints one JSON line (the same object bench.py puts under workloads.banded52269)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as g  # noqa: E402
import bench  # noqa: E402

pkg = g.load_package()
assert pkg.lib().cgb200_init(0) == 0, pkg.last_error()
peak, src = bench.measured_peak()
out = bench.long_row_leg(pkg, peak, src, 0)
tag = {k: v for k, v in os.environ.items() if k.startswith("CGB200_")}
print(json.dumps({"env": tag, "it_s": out["value"], "ms_per_step": out["ms_per_step"], "frac_iter": out["frac_of_peak"],
                  "k1_ms": out["roofline"]["avg_launch_ms"], "k1_frac": out["roofline"]["frac"],
                  "banded_format": out.get("banded_format")}))
