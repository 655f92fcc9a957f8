"""CPU, world_size 2 and 3 over gloo: the host-side logic of the multi-GPU layer.

The row-block plan (csrc/partition.cpp, reached through the C ABI cgb200_plan_partition) decides
halo widths, the local gather layout and who sends what.  Two real processes execute that plan with
gloo send/recv standing in for NCCL and the oracle's CPU operators standing in for the kernels, and
must reproduce the one-rank result: bit-exact for the matvec (row sums do not depend on the
partition) and to reduction-order tolerance for a CG solve (dot products are summed per rank).
The peer-memory layout (cgb200_plan_push: which rows each rank stores at which index of which
peer's gather vector) is executed the same way, one message per push item written at the planned
index, and must fill the gather vector exactly like the send/recv scheme; the single-reduction
recurrence runs on top of it against its CPU restatement."""
import os
import socket
import sys

import numpy as np
import pytest

import helpers as H

ROOT = H.ROOT


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _plan(pkg, world, rank, n, info, dense=0):
    import ctypes as C

    arr = (C.c_longlong * (4 * world))(*[int(v) for row in info for v in row])
    out = (C.c_longlong * 10)()
    assert pkg.lib().cgb200_plan_partition(world, rank, n, arr, dense, out) == 0, pkg.last_error()
    keys = ["allgather", "h_lo", "h_hi", "off", "ext_len", "col_shift", "send_lo", "send_hi", "recv_lo", "recv_hi"]
    return dict(zip(keys, [int(v) for v in out]))


def _push_plan(pkg, world, rank, n, info, dense=0):
    import ctypes as C

    arr = (C.c_longlong * (4 * world))(*[int(v) for row in info for v in row])
    out = (C.c_longlong * 42)()
    assert pkg.lib().cgb200_plan_push(world, rank, n, arr, dense, out) == 0, pkg.last_error()
    items = [dict(peer=int(out[1 + 4 * t]), begin=int(out[2 + 4 * t]), end=int(out[3 + 4 * t]),
                  dst=int(out[4 + 4 * t])) for t in range(int(out[0]))]
    waits = [int(out[34 + i]) for i in range(int(out[33]))]
    return items, waits


CUTS = {
    2: {"lap2d": [24 * 7], "lap3d": [8 * 7 * 6], "banded": [170], "dense": [40]},
    3: {"lap2d": [24 * 5, 24 * 11], "lap3d": [8 * 7 * 3, 8 * 7 * 7], "banded": [120, 290], "dense": [30, 61]},
}


def _worker(rank, world, port, case, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch
    import torch.distributed as dist

    import __graft_entry__ as g
    import helpers as HH

    pkg = g.load_package()
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        s = {"lap2d": lambda: HH.lap(24, 18), "lap3d": lambda: HH.lap(8, 7, 10), "banded": lambda: HH.banded(400, 9),
             "dense": lambda: HH.dense(96)}[case]()
        n = s.n
        # contiguous row blocks; uneven on purpose
        bounds = [0] + CUTS[world][case] + [n]
        row0 = bounds[rank]
        nloc = bounds[rank + 1] - bounds[rank]
        rp = s.row_ptr[row0:row0 + nloc + 1].astype(np.int64)
        lo_k, hi_k = int(rp[0]), int(rp[-1])
        cols_g = s.cols[lo_k:hi_k].copy()
        vals = s.vals[lo_k:hi_k].copy()
        rp_l = (rp - rp[0]).astype(np.int32)
        mine = torch.tensor([row0, nloc, int(cols_g.min()), int(cols_g.max())], dtype=torch.int64)
        allinfo = [torch.zeros(4, dtype=torch.int64) for _ in range(world)]
        dist.all_gather(allinfo, mine)
        info = [t.tolist() for t in allinfo]
        dense = 1 if case == "dense" else 0
        P = _plan(pkg, world, rank, n, info, dense)
        plans = [_plan(pkg, world, g_, n, info, dense) for g_ in range(world)]
        # all ranks must agree on the scheme and on matching message sizes
        assert all(pl["allgather"] == P["allgather"] for pl in plans)
        if not P["allgather"]:
            if rank == 0:
                assert P["h_lo"] == 0 and P["send_lo"] == 0
            else:
                lower = plans[rank - 1]
                assert P["send_lo"] == lower["recv_hi"] and P["recv_lo"] == lower["send_hi"]
            if rank == world - 1:
                assert P["h_hi"] == 0 and P["send_hi"] == 0
            else:
                upper = plans[rank + 1]
                assert P["send_hi"] == upper["recv_lo"] and P["recv_hi"] == upper["send_lo"]
            assert P["off"] % 16 == 0 and P["off"] >= P["h_lo"]
            assert P["ext_len"] == P["off"] + nloc + P["h_hi"]
        else:
            assert P["ext_len"] == n and P["off"] == row0 and P["col_shift"] == 0
        cols_l = (cols_g - P["col_shift"]).astype(np.int32)
        assert cols_l.min() >= 0 and cols_l.max() < P["ext_len"]

        O = HH.oracle()

        def exchange(p_ext):
            own = p_ext[P["off"]:P["off"] + nloc]
            if P["allgather"]:
                parts = [torch.zeros(info[g_][1], dtype=torch.float64) for g_ in range(world)]
                # gloo all_gather needs equal sizes: do it with broadcasts, as the NCCL path does
                for g_ in range(world):
                    t = torch.from_numpy(own.copy()) if g_ == rank else parts[g_]
                    dist.broadcast(t, g_)
                    p_ext[info[g_][0]:info[g_][0] + info[g_][1]] = t.numpy()
                return
            reqs = []
            if rank > 0:
                if P["send_lo"]:
                    reqs.append(dist.isend(torch.from_numpy(own[:P["send_lo"]].copy()), rank - 1))
                if P["recv_lo"]:
                    buf_lo = torch.zeros(P["recv_lo"], dtype=torch.float64)
                    reqs.append(dist.irecv(buf_lo, rank - 1))
            if rank + 1 < world:
                if P["send_hi"]:
                    reqs.append(dist.isend(torch.from_numpy(own[nloc - P["send_hi"]:].copy()), rank + 1))
                if P["recv_hi"]:
                    buf_hi = torch.zeros(P["recv_hi"], dtype=torch.float64)
                    reqs.append(dist.irecv(buf_hi, rank + 1))
            for r_ in reqs:
                r_.wait()
            if rank > 0 and P["recv_lo"]:
                p_ext[P["off"] - P["recv_lo"]:P["off"]] = buf_lo.numpy()
            if rank + 1 < world and P["recv_hi"]:
                p_ext[P["off"] + nloc:P["off"] + nloc + P["recv_hi"]] = buf_hi.numpy()

        # peer-memory layout: my push items, and (from the peers' plans) what lands where in mine
        my_items, my_waits = _push_plan(pkg, world, rank, n, info, dense)
        incoming = []
        for g_ in range(world):
            if g_ == rank:
                continue
            for it_ in _push_plan(pkg, world, g_, n, info, dense)[0]:
                if it_["peer"] == rank:
                    incoming.append((g_, it_))
        assert sorted(my_waits) == sorted({g_ for g_, _ in incoming})

        def exchange_push(p_ext):
            own = p_ext[P["off"]:P["off"] + nloc]
            reqs, bufs = [], []
            for it_ in my_items:
                reqs.append(dist.isend(torch.from_numpy(own[it_["begin"]:it_["end"]].copy()), it_["peer"]))
            for g_, it_ in incoming:
                buf = torch.zeros(it_["end"] - it_["begin"], dtype=torch.float64)
                bufs.append((it_["dst"], buf))
                reqs.append(dist.irecv(buf, g_))
            for r_ in reqs:
                r_.wait()
            for dst, buf in bufs:
                assert dst >= 0 and dst + len(buf) <= P["ext_len"]
                # a push never lands on the rows this rank owns
                assert dst + len(buf) <= P["off"] or dst >= P["off"] + nloc
                p_ext[dst:dst + len(buf)] = buf.numpy()

        def matvec(v_own, push=False):
            p_ext = np.zeros(P["ext_len"])
            p_ext[P["off"]:P["off"] + nloc] = v_own
            (exchange_push if push else exchange)(p_ext)
            out = np.empty(nloc)
            O.oracle_spmv(nloc, HH.iptr(rp_l), HH.iptr(cols_l), HH.dptr(vals), HH.dptr(p_ext), HH.dptr(out))
            return out

        def gsum(v):
            t = torch.tensor([v], dtype=torch.float64)
            dist.all_reduce(t)
            return float(t.item())

        # 1) matvec: bit-exact against the one-rank oracle
        v = HH.rhs(n, 77)
        want = np.empty(n)
        O.oracle_spmv(n, HH.iptr(s.row_ptr), HH.iptr(s.cols), HH.dptr(s.vals), HH.dptr(v), HH.dptr(want))
        got = matvec(v[row0:row0 + nloc].copy())
        ok_mv = bool(np.array_equal(got.view(np.uint64), want[row0:row0 + nloc].view(np.uint64)))
        got2 = matvec(v[row0:row0 + nloc].copy(), push=True)
        ok_mv = ok_mv and bool(np.array_equal(got2.view(np.uint64), want[row0:row0 + nloc].view(np.uint64)))

        # 2) the loop of cg.c:104-135 on row blocks with global scalars
        b = s.b[row0:row0 + nloc].copy()
        x = np.zeros(nloc)
        r = b.copy()
        p = b.copy()
        rr = gsum(float(np.dot(r, r)))
        M = 25 if case != "dense" else 4
        for k in range(M + 1):
            sv = matvec(p)
            alpha = rr / gsum(float(np.dot(p, sv)))
            x = x + alpha * p
            r = r - alpha * sv
            rr_new = gsum(float(np.dot(r, r)))
            if k == M:
                break
            p = r + (rr_new / rr) * p
            rr = rr_new
        xo, _, _, _ = HH.oracle_cg(s, M)
        diff = float(np.linalg.norm(x - xo[row0:row0 + nloc]) / np.linalg.norm(xo))

        # 3) the single-reduction recurrence (include/cgb200.h) over the peer-memory layout: one
        #    fused pair of sums per iteration, against its CPU restatement
        x = np.zeros(nloc)
        g_v = b.copy()
        d = np.zeros(nloc)
        sv = np.zeros(nloc)
        w = matvec(g_v, push=True)

        def gsum2(a_, b_):
            t = torch.tensor([a_, b_], dtype=torch.float64)
            dist.all_reduce(t)
            return float(t[0]), float(t[1])

        rr, rw = gsum2(float(np.dot(g_v, g_v)), float(np.dot(g_v, w)))
        alpha, beta = rr / rw, 0.0
        for k in range(M + 1):
            d = g_v + beta * d
            sv = w + beta * sv
            x = x + alpha * d
            g_v = g_v - alpha * sv
            w = matvec(g_v, push=True)
            rr_new, rw = gsum2(float(np.dot(g_v, g_v)), float(np.dot(g_v, w)))
            if k == M:
                break
            beta = rr_new / rr
            alpha = rr_new / (rw - beta * rr_new / alpha)
            rr = rr_new
        xs, its, _, _ = HH.oracle_cg_sr(s, M)
        assert its == M + 1
        diff_sr = float(np.linalg.norm(x - xs[row0:row0 + nloc]) / np.linalg.norm(xs))
        q.put((rank, ok_mv, max(diff, diff_sr), P))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,case", [(2, "lap2d"), (2, "lap3d"), (2, "banded"), (2, "dense"),
                                        (3, "lap3d"), (3, "banded"), (3, "dense")])
def test_row_blocks_reproduce_one_rank(world, case):
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, case, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=180) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, ok_mv, diff, P in res:
        assert ok_mv, f"rank {rank}: partitioned matvec differs from the one-rank oracle"
        assert diff <= 1e-8, (rank, diff)
        assert P["allgather"] == (1 if case == "dense" else 0)


def test_plan_rejects_gaps_and_handles_single_rank(pkg):
    import ctypes as C

    one = _plan(pkg, 1, 0, 100, [[0, 100, 0, 99]])
    assert one["allgather"] == 1 and one["off"] == 0 and one["ext_len"] == 100 and one["col_shift"] == 0
    arr = (C.c_longlong * 8)(0, 40, 0, 50, 50, 50, 30, 99)     # rows 40..49 owned by nobody
    out = (C.c_longlong * 10)()
    assert pkg.lib().cgb200_plan_partition(2, 0, 100, arr, 0, out) == -1
    # a halo wider than the neighbour's block forces the all-gather scheme on every rank
    wide = [[0, 10, 0, 25], [10, 10, 0, 39], [20, 20, 5, 39]]
    assert all(_plan(pkg, 3, r, 40, wide)["allgather"] == 1 for r in range(3))
    # 8 z-slabs of a 7-point stencil: one plane of halo each side, interior ranks symmetric
    pl, nzl = 64 * 64, 8
    info = [[g * pl * nzl, pl * nzl, max(0, g * pl * nzl - pl), min(8 * pl * nzl - 1, (g + 1) * pl * nzl - 1 + pl)] for g in range(8)]
    for g in range(8):
        P = _plan(pkg, 8, g, 8 * pl * nzl, info)
        assert P["allgather"] == 0
        assert P["h_lo"] == (pl if g > 0 else 0) and P["h_hi"] == (pl if g < 7 else 0)
        assert P["send_lo"] == (pl if g > 0 else 0) and P["send_hi"] == (pl if g < 7 else 0)
