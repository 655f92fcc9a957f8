"""Phase trace of the dense path's one-kernel iteration (k_cg_dense_iter), 1 or N ranks:
    python tools/dense_trace.py [n]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/dense_trace.py [n]
Prints, per rank, the average SM cycles from the start of the closing CTA to the end of each phase
(cgb200_system_trace) next to the device time per iteration.  Diagnostics only."""
import ctypes as C
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as g

rank, world, lr = (int(os.environ.get(k, d)) for k, d in (("RANK", 0), ("WORLD_SIZE", 1), ("LOCAL_RANK", 0)))
torch.cuda.set_device(lr)
pkg = g.load_package()
L = pkg.lib()
HAVE_TRACE = hasattr(L, "cgb200_system_trace")
if HAVE_TRACE:
    L.cgb200_system_trace.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_ulonglong)]
assert L.cgb200_init(lr) == 0, pkg.last_error()
if world > 1:
    import torch.distributed as dist

    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    idbuf = (C.c_ubyte * 128)()
    if rank == 0:
        assert L.cgb200_dist_unique_id(idbuf) == 0, pkg.last_error()
    t = torch.tensor(list(bytes(idbuf)), dtype=torch.uint8, device="cuda")
    dist.broadcast(t, 0)
    idbuf = (C.c_ubyte * 128).from_buffer_copy(bytes(t.cpu().tolist()))
    assert L.cgb200_dist_init(rank, world, idbuf) == 0, pkg.last_error()

n = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
per = n // world
d = pkg.DeviceSystem.dense(n, rank * per, per if rank < world - 1 else n - rank * per)
J, SOLVES = 5, 60


d.grid_hint = lambda: int(os.environ.get('GRID_HINT', '592'))


def run():
    tot, its = 0.0, 0
    for k in range(SOLVES):
        d.gen_rhs(100 + k)
        r = d.solve(J - 1, check_every=J)
        tot += r.seconds
        its += r.iterations
    return tot / its * 1e6


run()
us_plain = min(run(), run())
line = f"rank {rank}/{world} n={n} lib {os.environ.get('CGB200_LIBRARY', 'default')} it_us plain {us_plain:.1f}"
if HAVE_TRACE and d.one_kernel_iteration() and not os.environ.get("NO_TRACE"):
    assert L.cgb200_system_trace(d.h, 1, None) == 0, pkg.last_error()
    run()
    assert L.cgb200_system_trace(d.h, 1, None) == 0
    us_traced = run()
    out = (C.c_ulonglong * (16 + 2048))()
    assert L.cgb200_system_trace(d.h, 0, out) == 0
    cnt = max(out[7], 1)
    names = ["halo_wait", "gemv_done", "fixup_done", "alpha_known", "xr_done", "beta_known", "p_pushed"]
    avg = {nm: out[i] / cnt for i, nm in enumerate(names)}
    line += f" traced {us_traced:.1f} | cycles: " + " ".join(f"{k}={v:.0f}" for k, v in avg.items() if k != "fixup_done")
    line += (f" | ns: gemv_first_start_to_last_end={out[2] / cnt:.0f} start_spread={out[13] / cnt:.0f} "
             f"end_spread={out[14] / cnt:.0f} first_start_to_first_end={out[15] / cnt:.0f} "
             f"mean_cta_gemv={out[12] / cnt / max(d.grid_hint(), 1):.0f}")
    if rank == 0 and os.environ.get("TRACE_DUMP"):
        import json

        G = d.grid_hint()
        per = [(c_, int(out[16 + 1024 + c_]), out[16 + c_] / cnt) for c_ in range(G)]
        with open(os.environ["TRACE_DUMP"], "w") as f:
            json.dump({"n": n, "world": world, "launches": int(cnt), "per_cta": per}, f)
if world > 1:
    for r_ in range(world):
        if r_ == rank:
            print(line, flush=True)
        dist.barrier()
else:
    print(line, flush=True)
d.free()
if world > 1:
    L.cgb200_dist_finalize()
    dist.destroy_process_group()
