"""K2 + K3 as one launch (k_update_xrp) against the three-kernel loop on small systems: it/s and launches.
python tools/exp/fuse_vec_ab.py   (one GPU)"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import __graft_entry__ as g
import bench

pkg = g.load_package()
L = pkg.lib()
assert L.cgb200_init(-1) == 0


def banded():
    rp, cols, vals, b = bench.banded_host_system(52269, 172)
    return pkg.DeviceSystem.from_csr(52269, 0, rp, cols, vals, b)


for name, make, M in (("banded52269", banded, 400), ("lap2d_32x32", lambda: pkg.DeviceSystem.lap(32, 32, 1), 2000),
                      ("lap2d_512x512", lambda: pkg.DeviceSystem.lap(512, 512, 1), 1000),
                      ("lap3d_80", lambda: pkg.DeviceSystem.lap(80, 80, 80), 1000)):
    for fused in (0, 1, 0, 1):
        L.cgb200_set_fused_vector_update(fused)
        d = make()
        d.solve(50)
        best = 0.0
        for _ in range(3):
            d.gen_rhs(5) if name != "banded52269" else None
            r = d.solve(M - 1, check_every=M)
            best = max(best, r.iterations / r.seconds)
        print(f"{name:14s} fused={fused} it/s {best:10.0f}  us/it {1e6 / best:7.2f}  launches {r.kernel_launches} for {r.iterations} iterations", flush=True)
        d.free()
