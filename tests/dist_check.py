"""Multi-rank parity check, launched as
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tests/dist_check.py
Every rank builds its row block through the C ABI, the ranks solve together (NCCL), rank 0 gathers x
and compares with the CPU oracle on the whole system (tests/helpers.py; checker only)."""
import ctypes as C
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import __graft_entry__ as g
import helpers as H

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
pkg = g.load_package()
L = pkg.lib()
assert L.cgb200_init(lr) == 0, pkg.last_error()
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
idbuf = (C.c_ubyte * 128)()
if rank == 0 and world > 1:
    assert L.cgb200_dist_unique_id(idbuf) == 0, pkg.last_error()
t = torch.tensor(list(bytes(idbuf)), dtype=torch.uint8, device="cuda")
dist.broadcast(t, 0)
idbuf = (C.c_ubyte * 128).from_buffer_copy(bytes(t.cpu().tolist()))
assert L.cgb200_dist_init(rank, world, idbuf) == 0, pkg.last_error()
SR = pkg.SINGLE_REDUCTION


def gather_x(d, n, counts):
    xl = torch.from_numpy(d.x()).cuda()
    parts = [torch.empty(c, dtype=torch.float64, device="cuda") for c in counts]
    # uneven sizes: one broadcast per rank
    for r_ in range(world):
        buf = xl if r_ == rank else parts[r_]
        dist.broadcast(buf, r_)
        parts[r_] = buf
    return torch.cat(parts).cpu().numpy()


def blocks(n, align=1):
    base = (n // world) // align * align
    cnt = [base] * world
    cnt[-1] = n - base * (world - 1)
    return cnt


ok = True


def report(name, cond, detail):
    global ok
    ok = ok and bool(cond)
    if rank == 0:
        print(("PASS " if cond else "FAIL ") + name + " " + detail, flush=True)


# 1) device-generated 3D Laplacian, z-slabs, fixed count vs oracle; halo overlap path
nx, ny, nz = 48, 40, 8 * world
s = H.lap(nx, ny, nz)
nzl = nz // world
d = pkg.DeviceSystem.lap(nx, ny, nz, rank * nzl, nzl)
r = d.solve(60)
x = gather_x(d, s.n, [nx * ny * nzl] * world)
xo, it, _, _ = H.oracle_cg(s, 60)
rel = np.linalg.norm(x - xo) / np.linalg.norm(xo)
mode = "p2p" if L.cgb200_system_comm_is_p2p(d.h) else L.cgb200_dist_comm_mode().decode()
if rank == 0:
    print(f"comm {mode} (requested {os.environ.get('CGB200_COMM', 'p2p')}), world {world}", flush=True)
if world > 1 and os.environ.get("CGB200_COMM", "p2p") == "p2p" and os.environ.get("CGB200_REQUIRE_P2P"):
    assert mode == "p2p", "peer-memory mode was requested but the library fell back to NCCL"
report("lap3d fixed-count", r.iterations == it and rel <= 1e-8, f"iters {r.iterations}/{it} rel {rel:.2e} world {r.world}")
r = d.solve(60, variant=SR)
x = gather_x(d, s.n, [nx * ny * nzl] * world)
xs, its, _, _ = H.oracle_cg_sr(s, 60)
rel = np.linalg.norm(x - xs) / np.linalg.norm(xs)
report("lap3d fixed-count single-reduction", r.iterations == its and rel <= 1e-8, f"iters {r.iterations}/{its} rel {rel:.2e}")
# back to back solves on the same system (mailbox epochs carry over), alternating variants
for rep in range(3):
    r = d.solve(17, variant=rep % 2)
    x = gather_x(d, s.n, [nx * ny * nzl] * world)
    xo17, it17, _, _ = H.oracle_cg(s, 17)
    rel = np.linalg.norm(x - xo17) / np.linalg.norm(xo17)
    report(f"lap3d repeat {rep}", r.iterations == it17 and rel <= 1e-8, f"rel {rel:.2e}")
r = d.solve(100000, tol=1e-10)
xo, it, _, st = H.oracle_cg(s, 100000, tol=1e-10, guard=1)
x = gather_x(d, s.n, [nx * ny * nzl] * world)
tr = d.true_relres()
rel = np.linalg.norm(x - xo) / np.linalg.norm(xo)
report("lap3d tolerance", abs(r.iterations - it) <= 2 and tr <= 1.05e-10 and rel <= 1e-8,
       f"iters {r.iterations}/{it} true relres {tr:.2e} rel {rel:.2e}")
r = d.solve(100000, tol=1e-10, variant=SR)
x = gather_x(d, s.n, [nx * ny * nzl] * world)
tr = d.true_relres()
rel = np.linalg.norm(x - xo) / np.linalg.norm(xo)
report("lap3d tolerance single-reduction", abs(r.iterations - it) <= 2 and tr <= 1.05e-10 and rel <= 1e-8,
       f"iters {r.iterations}/{it} true relres {tr:.2e} rel {rel:.2e}")
# solve options beyond the reference: initial guess + Jacobi preconditioner through the exchange
if hasattr(L, "cgb200_system_set_x"):
    nl_ = nx * ny * nzl
    x0 = np.random.default_rng(3).standard_normal(s.n)
    for jac, warm in ((True, False), (False, True), (True, True)):
        xo_, it_, _, _ = H.oracle_pcg(s, 25, x0=x0 if warm else None, jacobi=jac)
        r = d.solve(25, jacobi=jac, x0=x0[rank * nl_:(rank + 1) * nl_] if warm else None)
        x = gather_x(d, s.n, [nl_] * world)
        rel = np.linalg.norm(x - xo_) / np.linalg.norm(xo_)
        report(f"lap3d options jacobi={int(jac)} x0={int(warm)}", r.iterations == it_ and rel <= 1e-8, f"rel {rel:.2e}")
    r = d.solve(100000, tol=1e-10, jacobi=True, x0=True)
    tr = d.true_relres()
    report("lap3d options tolerance restart", r.status == pkg.CONVERGED and tr <= 1.05e-10, f"iters {r.iterations} true relres {tr:.2e}")
# plain products through the exchange (operator-API path), bit-exact per row
v = H.rhs(s.n, 77)
want = np.empty(s.n)
H.oracle().oracle_spmv(s.n, H.iptr(s.row_ptr), H.iptr(s.cols), H.dptr(s.vals), H.dptr(v), H.dptr(want))
nl_ = nx * ny * nzl
for rep in range(2):
    got = d.matvec(v[rank * nl_:(rank + 1) * nl_])
    okb = np.array_equal(got.view(np.uint64), want[rank * nl_:(rank + 1) * nl_].view(np.uint64))
    tt = torch.tensor([1.0 if okb else 0.0], device="cuda")
    dist.all_reduce(tt, op=dist.ReduceOp.MIN)
    report(f"lap3d plain matvec bit-exact {rep}", tt.item() == 1.0, "")
d.free()

# 2) bigger slab so interior and boundary tiles both exist (tile = 512 rows)
nx, ny, nz = 64, 64, 16 * world
nzl = nz // world
d = pkg.DeviceSystem.lap(nx, ny, nz, rank * nzl, nzl)
s = H.lap(nx, ny, nz)
r = d.solve(40)
x = gather_x(d, s.n, [nx * ny * nzl] * world)
xo, it, _, _ = H.oracle_cg(s, 40)
rel = np.linalg.norm(x - xo) / np.linalg.norm(xo)
report("lap3d split-launch", r.iterations == it and rel <= 1e-8, f"rel {rel:.2e}")
r = d.solve(40, variant=SR)
x = gather_x(d, s.n, [nx * ny * nzl] * world)
rel = np.linalg.norm(x - xo) / np.linalg.norm(xo)
report("lap3d split-launch single-reduction", r.iterations == it and rel <= 1e-8, f"rel {rel:.2e}")
# the same slab in the row-pattern dictionary format: interior + boundary tiles, halo waits
if hasattr(L, "cgb200_system_use_pattern_format"):
    okp = d.use_pattern_format(True)
    for variant, nm in ((0, "classic"), (SR, "single-reduction")):
        r = d.solve(40, variant=variant)
        x = gather_x(d, s.n, [nx * ny * nzl] * world)
        rel = np.linalg.norm(x - xo) / np.linalg.norm(xo)
        report(f"lap3d pattern format {nm}", okp and r.iterations == it and rel <= 1e-8, f"rel {rel:.2e}")
    v = H.rhs(s.n, 78)
    want = np.empty(s.n)
    H.oracle().oracle_spmv(s.n, H.iptr(s.row_ptr), H.iptr(s.cols), H.dptr(s.vals), H.dptr(v), H.dptr(want))
    nl_ = nx * ny * nzl
    got = d.matvec(v[rank * nl_:(rank + 1) * nl_])
    okb = np.array_equal(got.view(np.uint64), want[rank * nl_:(rank + 1) * nl_].view(np.uint64))
    tt = torch.tensor([1.0 if okb else 0.0], device="cuda")
    dist.all_reduce(tt, op=dist.ReduceOp.MIN)
    report("lap3d pattern format plain matvec bit-exact", tt.item() == 1.0, "")
d.free()

# 3) dense, all-gather mode
n = 1024 * world
s = H.dense(n)
nl = n // world
d = pkg.DeviceSystem.dense(n, rank * nl, nl)
r = d.solve(200, tol=1e-10)
x = gather_x(d, n, [nl] * world)
xo, it, _, _ = H.oracle_cg(s, 200, tol=1e-10, guard=1)
rel = np.linalg.norm(x - xo) / np.linalg.norm(xo)
report("dense all-gather", abs(r.iterations - it) <= 2 and rel <= 1e-8, f"iters {r.iterations}/{it} rel {rel:.2e}")
r = d.solve(200, tol=1e-10, variant=SR)
x = gather_x(d, n, [nl] * world)
rel = np.linalg.norm(x - xo) / np.linalg.norm(xo)
report("dense all-gather single-reduction", abs(r.iterations - it) <= 2 and rel <= 1e-8, f"iters {r.iterations}/{it} rel {rel:.2e}")
d.free()

# 3b) a dense system big enough for the matvec's dynamic tail (>= 64 MB per rank), in uneven row
#     blocks that are not multiples of the row-group height and only 16-byte aligned
n = 6144 + 1024 * world
s = H.dense(n)
cnt = blocks(n)
if world > 1:
    cnt[0] += 2
    cnt[-1] -= 2
row0 = sum(cnt[:rank])
d = pkg.DeviceSystem.dense(n, row0, cnt[rank])
xo, it, _, _ = H.oracle_cg(s, 4)
for variant, nm in ((0, "classic (one-kernel iteration)"), (SR, "single-reduction")):
    r = d.solve(4, variant=variant)
    x = gather_x(d, n, cnt)
    rel = np.linalg.norm(x - xo) / np.linalg.norm(xo)
    report(f"dense uneven blocks, dynamic tail, {nm}", r.iterations == it and rel <= 1e-8,
           f"iters {r.iterations}/{it} rel {rel:.2e} one-kernel {d.one_kernel_iteration()}")
if hasattr(L, "cgb200_system_set_x"):
    x0 = np.random.default_rng(4).standard_normal(n)
    xo_, it_, _, _ = H.oracle_pcg(s, 4, x0=x0, jacobi=True)
    r = d.solve(4, jacobi=True, x0=x0[row0:row0 + cnt[rank]])
    x = gather_x(d, n, cnt)
    rel = np.linalg.norm(x - xo_) / np.linalg.norm(xo_)
    report("dense uneven blocks, jacobi + x0", r.iterations == it_ and rel <= 1e-8, f"rel {rel:.2e}")
x1 = None
for rep in range(2):   # the dynamic tail must not cost reproducibility
    d.gen_rhs(5)
    d.solve(4)
    xr = d.x()
    if x1 is None:
        x1 = xr
same = np.array_equal(x1.view(np.uint64), xr.view(np.uint64))
tt = torch.tensor([1.0 if same else 0.0], device="cuda")
dist.all_reduce(tt, op=dist.ReduceOp.MIN)
report("dense dynamic tail run-to-run bits", tt.item() == 1.0, "")
d.free()

# 4) host CSR row blocks, uneven, banded (neighbour mode) and wide band (all-gather fallback)
# (hb = 60: 121 entries per row, the long-row matvec kernel behind a neighbour halo; hb = 900 the same
# kernel in all-gather mode)
for hb, name in ((7, "banded neighbour"), (60, "banded neighbour long rows"), (900, "banded wide")):
    s = H.banded(3001, hb)
    cnt = blocks(s.n)
    row0 = sum(cnt[:rank])
    nl = cnt[rank]
    rp = s.row_ptr[row0:row0 + nl + 1]
    d = pkg.DeviceSystem.from_csr(s.n, row0, rp - rp[0], s.cols[rp[0]:rp[-1]], s.vals[rp[0]:rp[-1]], s.b[row0:row0 + nl])
    for variant in (0, SR):
        r = d.solve(15, variant=variant)
        x = gather_x(d, s.n, cnt)
        xo, it, _, _ = H.oracle_cg(s, 15)
        rel = np.linalg.norm(x - xo) / np.linalg.norm(xo)
        report(f"from_csr {name} variant {variant}", r.iterations == it and rel <= 1e-8, f"rel {rel:.2e}")
    if hasattr(L, "cgb200_system_set_x"):
        x0 = np.random.default_rng(5).standard_normal(s.n)
        xo_, it_, _, _ = H.oracle_pcg(s, 15, x0=x0, jacobi=True)
        r = d.solve(15, jacobi=True, x0=x0[row0:row0 + nl])
        x = gather_x(d, s.n, cnt)
        rel = np.linalg.norm(x - xo_) / np.linalg.norm(xo_)
        report(f"from_csr {name} jacobi + x0", r.iterations == it_ and rel <= 1e-8, f"rel {rel:.2e}")
    d.free()

# 5) run under CGB200_GUARD=1: no kernel stored outside a device buffer on any rank (csrc/devmem.h)
if hasattr(L, "cgb200_guard_check"):
    L.cgb200_guard_check.restype = C.c_longlong
    bad = int(L.cgb200_guard_check())
    if bad >= 0:
        tg = torch.tensor([float(bad)], dtype=torch.float64, device="cuda")
        dist.all_reduce(tg, op=dist.ReduceOp.SUM)
        report("guard bands intact on every rank", tg.item() == 0.0, f"{int(tg.item())} damaged")

L.cgb200_dist_finalize()
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
