"""Dense matvec time against the number of rows of a 32768-column row block on ONE GPU: separates
the fixed cost of a launch (ramp, reductions, barriers) from the streaming rate.
usage: gemv_rows_sweep.py [ncols]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as g

pkg = g.load_package()
L = pkg.lib()
assert L.cgb200_init(0) == 0, pkg.last_error()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
rows_list = [512, 1024, 2048, 4096, 8192, 16384, 32768]
res = []
for rows in rows_list:
    if rows > n:
        continue
    d = pkg.DeviceSystem.dense(n, 0, rows)
    t1 = d.time_kernel(1, 40)
    t6 = d.time_kernel(6, 40) if d.one_kernel_iteration() else float("nan")
    gb = 8.0 * rows * n / 1e9
    res.append((rows, t1, t6))
    print(f"rows {rows:6d}  bytes {gb:7.3f} GB  K1 {t1*1e3:8.1f} us {gb/t1*1e3:7.0f} GB/s   one-kernel {t6*1e3:8.1f} us {gb/t6*1e3:7.0f} GB/s", flush=True)
    d.free()
if len(res) >= 2:
    (r0, a0, b0), (r1, a1, b1) = res[-3], res[-1]
    for nm, x0, x1 in (("K1", a0, a1), ("one-kernel", b0, b1)):
        slope = (x1 - x0) / (r1 - r0)
        print(f"{nm}: fixed {1e3*(x0 - slope*r0):.1f} us, streaming {8.0*n/slope/1e6:.0f} GB/s (fit through rows {r0} and {r1})")
