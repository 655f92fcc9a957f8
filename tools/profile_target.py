"""Small fixed workload for ncu captures: a few CG iterations of one BASELINE config on 1 GPU.
usage: profile_target.py {dense|lap2d|lap3d512|lap3d384|banded|bandeddia} [iterations] [classic|sr] [pattern]
(banded: the course-sized long-row system of bench.py's `banded52269` line, n = 52269, 345 per row;
bandeddia: the same system in the optional diagonal storage, csrc/banded.cu)"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as g

pkg = g.load_package()
L = pkg.lib()
assert L.cgb200_init(-1) == 0, pkg.last_error()
w = sys.argv[1]
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 8


def _banded():
    import bench

    rp, cols, vals, b = bench.banded_host_system(52269, 172)
    return pkg.DeviceSystem.from_csr(52269, 0, rp, cols, vals, b)


d = {
    "banded": _banded,
    "bandeddia": _banded,
    "dense": lambda: pkg.DeviceSystem.dense(32768),
    "lap2d": lambda: pkg.DeviceSystem.lap(4096, 4096, 1),
    "lap3d512": lambda: pkg.DeviceSystem.lap(512, 512, 512),
    "lap3d384": lambda: pkg.DeviceSystem.lap(384, 384, 384),
}[w]()
variant = pkg.SINGLE_REDUCTION if (len(sys.argv) > 3 and sys.argv[3] == "sr") else pkg.CLASSIC
if w == "bandeddia":
    assert d.use_banded_format(True), pkg.last_error()
if len(sys.argv) > 4 and sys.argv[4] == "pattern":
    assert d.use_pattern_format(True), pkg.last_error()
r = d.solve(iters - 1, check_every=iters, variant=variant)
print(w, "iterations", r.iterations, "status", r.status, "relres", r.relres, "ms/iter", r.seconds * 1e3 / max(r.iterations, 1))
d.free()
