"""Times the multi-rank form of the long-row matvec (k_spmv_rowblock<.., COHERENT>: every CTA behind the
peers' halo flags, gathers at L2) on the course-sized banded system cut into row blocks:
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port P tools/longrow_multi.py
Prints it/s of 300 fixed-count iterations and the matvec kernel's time, per exchange mode in use."""
import ctypes as C
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as g  # noqa: E402
import bench  # noqa: E402

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
pkg = g.load_package()
L = pkg.lib()
assert L.cgb200_init(lr) == 0, pkg.last_error()
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
idbuf = (C.c_ubyte * 128)()
if rank == 0:
    assert L.cgb200_dist_unique_id(idbuf) == 0, pkg.last_error()
t = torch.tensor(list(bytes(idbuf)), dtype=torch.uint8, device="cuda")
dist.broadcast(t, 0)
idbuf = (C.c_ubyte * 128).from_buffer_copy(bytes(t.cpu().tolist()))
assert L.cgb200_dist_init(rank, world, idbuf) == 0, pkg.last_error()

n, hb = 52269, 172
rp, cols, vals, b = bench.banded_host_system(n, hb)
r0, r1 = n * rank // world, n * (rank + 1) // world
k0, k1 = rp[r0], rp[r1]
d = pkg.DeviceSystem.from_csr(n, r0, rp[r0:r1 + 1] - k0, cols[k0:k1], vals[k0:k1], b[r0:r1])
mode = "p2p" if L.cgb200_system_comm_is_p2p(d.h) else "nccl"
d.solve(20)
best = 0.0
for _ in range(3):
    r = d.solve(299)
    best = max(best, r.iterations / r.seconds)
k1ms = d.time_kernel(1, 30)
ib, k1b = d.bytes()
tt = torch.tensor([best], dtype=torch.float64, device="cuda")
dist.all_reduce(tt, op=dist.ReduceOp.MIN)
if rank == 0:
    print(f"banded52269 world {world} comm {mode}: {tt.item():.0f} it/s ({1e6 / tt.item():.1f} us per iteration); "
          f"matvec alone on this rank's block {k1ms * 1e3:.1f} us = {k1b / k1ms / 1e6:.0f} GB/s", flush=True)
d.free()
L.cgb200_dist_finalize()
dist.destroy_process_group()
