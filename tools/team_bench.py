"""conj_grad / cgb200_solve on HOST structs with CGB200_GPUS = 1, 2, ... (csrc/team.cu): what an
unmodified caller of the reference's entry point gets from several GPUs.

    python tools/team_bench.py [--gpus 1,2] [--dense 32768] [--lap 256]

One child process per GPU count (the team is started once per process).  Each child builds the
system on the host with the oracle's generators (the same arrays at every count), calls
conj_grad(4, A, b, &x) cold (upload + cut into row blocks), then times warm solves with b refreshed
in place -- every solve copies b to and x from every GPU.  x after the cold call is compared across
GPU counts.  Wall clock around each C call, summed; one JSON line per (count, system).
"""
import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def child(args):
    import ctypes as C

    import numpy as np

    sys.path.insert(0, os.path.join(ROOT, "tests"))
    sys.path.insert(0, ROOT)
    import __graft_entry__ as g
    import helpers as H

    pkg = g.load_package()
    L = pkg.lib()
    systems = []
    if args.dense:
        systems.append(("dense%d" % args.dense, lambda: H.dense(args.dense), 200, 1e-10))
    if args.lap:
        systems.append(("lap3d%d" % args.lap, lambda: H.lap(args.lap, args.lap, args.lap), 300, 0.0))
    for name, make, max_iter, tol in systems:
        t = time.perf_counter()
        s = make()
        gen_s = time.perf_counter() - t
        A = pkg.MvSparse(s.n, s.nnz, H.dptr(s.vals), H.iptr(s.cols), H.iptr(s.row_ptr))
        b = pkg.MvSparse(s.n, s.n, H.dptr(s.b), None, None)
        xp = pkg.MvSparseP()
        t = time.perf_counter()
        assert L.conj_grad(4, C.byref(A), C.byref(b), C.byref(xp)) == 0, pkg.last_error()
        cold_s = time.perf_counter() - t
        x = np.ctypeslib.as_array(xp.contents.values, shape=(s.n,)).copy()
        np.save(os.path.join(args.tmp, f"x_{name}_{args.child}.npy"), x)
        gpus = int(L.cgb200_gpus_in_use())
        o = pkg.Opts(max_iter, tol, 0, 0, 0)
        res = pkg.Result()
        rng = np.random.default_rng(3)
        iters = 0
        solves = 6
        b0 = s.b.copy()
        st0 = pkg.cache_stats()
        warm_s = 0.0
        for _ in range(solves):
            s.b[:] = rng.random(s.n) - 0.5     # (the caller's own work: outside the clock)
            xq = pkg.MvSparseP()
            t = time.perf_counter()
            assert L.cgb200_solve(C.byref(A), C.byref(b), C.byref(xq), C.byref(o), C.byref(res)) == 0, pkg.last_error()
            warm_s += time.perf_counter() - t
            iters += res.iterations
            L.cgb200_free_all(xq)
        st1 = pkg.cache_stats()
        s.b[:] = b0
        print("TEAM_BENCH " + json.dumps({
            "system": name, "n": int(s.n), "nnz": int(s.nnz), "gpus_requested": args.child, "gpus_in_use": gpus,
            "world_of_last_solve": int(res.world), "cold_conj_grad_4_s": round(cold_s, 3),
            "warm_it_s": round(iters / warm_s, 1), "iterations_per_solve": iters / solves,
            "ms_per_solve": round(warm_s / solves * 1e3, 3),
            "calls_settled_by_write_tracking": st1["tracked_clean"] - st0["tracked_clean"],
            "content_hashes": st1["hashes"] - st0["hashes"], "uploads": st1["uploads"] - st0["uploads"],
            "host_generation_s": round(gen_s, 1)}), flush=True)
        L.cgb200_invalidate(C.byref(A))
        L.cgb200_free_all(xp)
        del s
    L.cgb200_shutdown()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", default="1,2")
    ap.add_argument("--dense", type=int, default=32768)
    ap.add_argument("--lap", type=int, default=256)
    ap.add_argument("--child", type=int, default=0)
    ap.add_argument("--tmp", default="/tmp")
    args = ap.parse_args()
    if args.child:
        return child(args)
    import numpy as np

    counts = [int(c) for c in args.gpus.split(",")]
    rows = []
    for c in counts:
        out = subprocess.run([sys.executable, os.path.abspath(__file__), "--child", str(c), "--dense", str(args.dense),
                              "--lap", str(args.lap), "--tmp", args.tmp], env=dict(os.environ, CGB200_GPUS=str(c)),
                             stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=1500)
        if out.returncode != 0:
            print(out.stdout[-3000:])
            return 1
        for ln in out.stdout.splitlines():
            if ln.startswith("TEAM_BENCH "):
                rows.append(json.loads(ln[len("TEAM_BENCH "):]))
                print(ln, flush=True)
            elif ln.startswith("[cgb200]"):
                print(ln, flush=True)
    # x of the cold conj_grad(4) call: every count against the first
    for name in sorted({r["system"] for r in rows}):
        ref = np.load(os.path.join(args.tmp, f"x_{name}_{counts[0]}.npy"))
        for c in counts[1:]:
            x = np.load(os.path.join(args.tmp, f"x_{name}_{c}.npy"))
            print("TEAM_PARITY " + json.dumps({"system": name, "gpus": c, "against_gpus": counts[0],
                                               "rel_l2": float(np.linalg.norm(x - ref) / np.linalg.norm(ref))}), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
