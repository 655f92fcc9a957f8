"""Long-row matvec alone on banded systems of growing size (fixed cost vs streaming rate):
    python tools/longrow_scale.py   -> one line per n: matvec us, GB/s, fraction of the measured peak."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as g  # noqa: E402
import bench  # noqa: E402

pkg = g.load_package()
assert pkg.lib().cgb200_init(0) == 0, pkg.last_error()
peak, _ = bench.measured_peak()
for n in (13067, 26134, 52269, 104538, 209076):
    rp, cols, vals, b = bench.banded_host_system(n, 172)
    d = pkg.DeviceSystem.from_csr(n, 0, rp, cols, vals, b)
    _, k1b = d.bytes()
    t = min(d.time_kernel(1, 30) for _ in range(3))
    print(f"n {n:7d} nnz {len(vals):9d} bytes {k1b / 1e6:7.1f} MB  matvec {t * 1e3:7.2f} us  {k1b / t / 1e6:7.0f} GB/s  frac {k1b / t / 1e6 / peak:.3f}",
          flush=True)
    d.free()
