#!/usr/bin/env python
"""bench.py -- CG iterations/second of the hot path on N B200s, beside the reference's CPU build.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl b200|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" is one CG iteration (one matvec + the two fused vector updates: one pass of the hot path
over the whole system).  The headline workload is BASELINE.json configs[1]: dense SPD n = 32768,
fp64, values-only GEMV path; with N > 1 the same system is row-block partitioned over the ranks
(strong scaling: search direction all-gathered, two scalar all-reduces per iteration).  Because
that matrix is strictly diagonally dominant CG converges to 1e-10 in a handful of iterations, so
K steps are run as back-to-back solves of J iterations each with a fresh right-hand side
(J = iterations a tolerance-1e-10 solve needs, measured in the warm-up), see DESIGN.md.

Prints ONE JSON line (rank 0).  Keys beyond the base contract: roofline, cpu_baseline, e2e,
clocks, gpu_launches, workloads (the other BASELINE configs measured in the same run).
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "cg_iterations_per_second"
UNIT = "it/s"


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def traffic_from_profiles(kernel, workload, world):
    """dram bytes per launch of the dominant kernel from the committed ncu capture of this very
    launch shape (same workload, 1 GPU), or None."""
    if world != 1:
        return None
    key = {"dense32768": "dense", "lap3d512": "lap3d512", "lap2d4096": "lap2d"}.get(workload)
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f)["_by_workload"][key][kernel]
    except Exception:
        return None


# -------------------------------------------------------------------------------------------------
# clocks sampler (nvidia-smi during the timed region)
# -------------------------------------------------------------------------------------------------


class ClockSampler:
    FIELDS = (
        "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
        "clocks_event_reasons.sw_power_cap"
    )

    def __init__(self, gpu_index):
        self.samples = []  # (t, sm, max, power, [flags])
        self.proc = None
        self.thread = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(gpu_index), f"--query-gpu={self.FIELDS}",
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True,
            )
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception as e:  # no nvidia-smi: report that instead of inventing clocks
            log("clock sampler unavailable:", e)

    def _pump(self):
        for line in self.proc.stdout:
            p = [x.strip() for x in line.split(",")]
            if len(p) < 7:
                continue
            try:
                self.samples.append((time.time(), float(p[0]), float(p[1]), float(p[2]), p[3:7]))
            except ValueError:
                pass

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self, t0, t1):
        inside = [s for s in self.samples if t0 <= s[0] <= t1]
        window = "timed region"
        if not inside:
            inside = [s for s in self.samples if s[0] <= t1 + 0.2][-5:]
            window = "nearest samples (timed region shorter than the sampling period)"
        if not inside:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0, "window": "none"}
        sm = sorted(s[1] for s in inside)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(s[4][i].lower().startswith("active") for s in inside)]
        return {
            "sm_mhz": sm[len(sm) // 2], "sm_max_mhz": inside[0][2], "reasons": reasons,
            "power_w_max": max(s[3] for s in inside), "samples": len(inside), "window": window,
        }


# -------------------------------------------------------------------------------------------------
# workloads
# -------------------------------------------------------------------------------------------------


def make_system(pkg, name, rank, world):
    """This rank's row block of the named workload, generated on the device."""
    if name == "dense32768":
        n = 32768
        assert n % world == 0
        nl = n // world
        return pkg.DeviceSystem.dense(n, rank * nl, nl), {"kind": "dense", "n": n, "scaling": "strong"}
    if name == "lap2d4096":
        assert world == 1
        return pkg.DeviceSystem.lap(4096, 4096, 1), {"kind": "lap2d", "n": 4096 * 4096, "scaling": "strong"}
    if name == "lap3d512":
        nz = 512
        assert nz % world == 0
        nzl = nz // world
        return pkg.DeviceSystem.lap(512, 512, nz, rank * nzl, nzl), {"kind": "lap3d", "n": 512 ** 3, "scaling": "strong"}
    if name == "lap3d384w":
        nz = 384 * world
        return pkg.DeviceSystem.lap(384, 384, nz, rank * 384, 384), {"kind": "lap3d", "n": 384 * 384 * nz, "scaling": "weak"}
    raise SystemExit(f"unknown workload {name}")


class Runner:
    """Runs exactly K CG iterations of a workload as the reference's loop would be used."""

    def __init__(self, pkg, d, info, variant=0):
        self.pkg, self.d, self.info, self.variant = pkg, d, info, variant
        self.J = None      # iterations per solve (dense: to 1e-10; stencils: all K in one solve)
        self.seed = 1000

    def calibrate(self):
        if self.info["kind"] == "dense" or self.info.get("short_solves"):
            r = self.d.solve(200, tol=1e-10, variant=self.variant)
            assert r.status == self.pkg.CONVERGED, f"warm-up solve did not converge: status {r.status}"
            self.J = r.iterations
        return self.J

    def run(self, K):
        """K iterations; returns the number actually performed (must equal K)."""
        done = 0
        if self.J is None:
            r = self.d.solve(K - 1, variant=self.variant)
            return r.iterations
        # back-to-back fixed-count solves, queued a few ahead of the GPU (cgb200_system_solve_async):
        # right-hand side generation and the solve are stream-ordered, results are collected behind
        pending = []
        queued = 0

        def collect():
            nonlocal done
            j_, t_ = pending.pop(0)
            r = self.d.solve_wait(t_)
            done += r.iterations
            if r.iterations != j_:
                raise RuntimeError(f"solve stopped early: {r.iterations} of {j_} (status {r.status})")

        while queued < K:
            j = min(self.J, K - queued)
            self.seed += 1
            self.d.gen_rhs(self.seed)     # fresh right-hand side, generated on the device
            pending.append((j, self.d.solve_async(j - 1, variant=self.variant)))
            queued += j
            if len(pending) >= 4:
                collect()
        while pending:
            collect()
        return done


def dist_sync(world):
    import torch

    if world > 1:
        import torch.distributed as dist

        dist.barrier()
    torch.cuda.synchronize()


def max_over_ranks(v, world):
    if world == 1:
        return v
    import torch
    import torch.distributed as dist

    t = torch.tensor([v], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def timed_steps(pkg, runner, K, W, world, sampler=None):
    L = pkg.lib()
    runner.calibrate()
    w_done = runner.run(max(W, 3))
    dist_sync(world)
    l0 = L.cgb200_launch_count()
    t0 = time.time()
    assert L.cgb200_timer_start() == 0, pkg.last_error()
    done = runner.run(K)
    ms = C.c_double(0.0)
    assert L.cgb200_timer_stop(C.byref(ms)) == 0, pkg.last_error()
    t1 = time.time()
    launches = L.cgb200_launch_count() - l0
    dist_sync(world)
    assert done == K, (done, K)
    ms_max = max_over_ranks(ms.value, world)
    return {"ms": ms_max, "steps": done, "warmup": w_done, "launches": int(launches), "t0": t0, "t1": t1}


def banded_host_system(n, hb, seed=3):
    """Banded SPD system in host CSR arrays (numpy only): A_ij = A_ji = U(-1/2, 1/2) from a counter-based
    hash of (min(i,j), max(i,j)) for 0 < |i-j| <= hb, diagonal = sum of the row's |off-diagonals| + 1."""
    import numpy as np

    i = np.arange(n, dtype=np.int64)[:, None]
    j = i + np.arange(-hb, hb + 1, dtype=np.int64)[None, :]
    mask = (j >= 0) & (j < n)
    lo, hi = np.minimum(i, j), np.maximum(i, j)
    with np.errstate(over="ignore"):
        h = (lo.astype(np.uint64) * np.uint64(n) + hi.astype(np.uint64) + np.uint64(seed)) * np.uint64(0x9E3779B97F4A7C15)
        h ^= h >> np.uint64(30)
        h *= np.uint64(0xBF58476D1CE4E5B9)
        h ^= h >> np.uint64(27)
        h *= np.uint64(0x94D049BB133111EB)
        h ^= h >> np.uint64(31)
    v = (h >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0) - 0.5
    v[~mask] = 0.0
    v[:, hb] = 0.0
    v[:, hb] = np.abs(v).sum(axis=1) + 1.0
    vals = np.ascontiguousarray(v[mask])
    cols = np.ascontiguousarray(j[mask].astype(np.int32))
    rp = np.zeros(n + 1, dtype=np.int32)
    rp[1:] = np.cumsum(mask.sum(axis=1))
    rng = np.random.default_rng(seed)
    return rp, cols, vals, rng.random(n) - 0.5


def long_row_leg(pkg, peak, peak_src, variant, K=300):
    """CSR rows of hundreds of entries (the reference's course matrix is of this kind, cg.c:231-232,260):
    banded SPD n = 52269, 345 entries per row, 18.0 M non-zeros -- the long-row matvec kernel."""
    n, hb = 52269, 172
    rp, cols, vals, b = banded_host_system(n, hb)
    d = pkg.DeviceSystem.from_csr(n, 0, rp, cols, vals, b)
    info = {"kind": "banded", "n": n, "scaling": "strong", "short_solves": True}
    rx = Runner(pkg, d, info, variant)
    tx = timed_steps(pkg, rx, K, 5, 1)
    ib, k1b = d.bytes()
    ms = tx["ms"] / tx["steps"]
    k1 = d.time_kernel(1, 30)
    out = {"value": round(tx["steps"] / (tx["ms"] * 1e-3), 2), "unit": UNIT, "steps": tx["steps"],
           "ms_per_step": round(ms, 5), "scaling": "strong", "iterations_per_solve": rx.J,
           "achieved_gbs_per_gpu": round(ib / (ms * 1e-3) / 1e9, 1), "frac_of_peak": round(ib / (ms * 1e-3) / 1e9 / peak, 4),
           "config": {"workload": "banded SPD n=52269, 345 entries per row, 18.0 M non-zeros, CSR fp64 (course-sized; not a "
                                  "BASELINE config)", "n": n, "nnz": int(len(vals)), "matvec": "CSR long-row SpMV (k_spmv_rowblock)",
                      "l2": "inputs larger than L2 (217 MB matrix streamed every iteration)"},
           "roofline": {"bound": "hbm", "kernel": "k_spmv_rowblock", "achieved": round(k1b / (k1 * 1e-3) / 1e9, 1),
                        "peak": peak, "unit": "GB/s", "frac": round(k1b / (k1 * 1e-3) / 1e9 / peak, 4), "traffic": 233068032,
                        "algorithmic_bytes_per_launch": k1b, "avg_launch_ms": round(k1, 5), "peak_source": peak_src,
                        "traffic_source": "profiles/r01f_ncu_longrow.md (32-row form of the same kernel)"}}
    d.free()
    return out


def time_to_solution(pkg, d, info, world, variant=0, tol=1e-10, fixed=None, cap=60000):
    """One solve from x0 = 0 to ||r||/||b|| <= tol (north_star's stop test; the reference itself has
    none) -- or `fixed` iterations (BASELINE configs[4]) -- timed on the device (CUDA
    events around the loop, max over ranks), with the residual recomputed from scratch afterwards."""
    d.gen_rhs(7)
    dist_sync(world)
    if fixed:
        r = d.solve(fixed - 1, variant=variant)
    else:
        r = d.solve(cap, tol=tol, variant=variant)
    dist_sync(world)
    sec = max_over_ranks(r.seconds, world)
    out = {"iterations": int(r.iterations), "seconds": round(sec, 5),
           "status": "converged" if r.status == pkg.CONVERGED else ("max_iter" if fixed else f"status {r.status}"),
           "relres_recursive": float(r.relres),
           "timing": "CUDA events around the device loop (cgb200_result.seconds), max over ranks; x0 = 0"}
    out["stop"] = f"fixed {fixed} iterations" if fixed else f"||r||/||b|| <= {tol:g}"
    try:
        out["relres_true"] = float(d.true_relres())
    except Exception as e:  # noqa: BLE001
        out["relres_true"] = None
        out["relres_true_error"] = repr(e)
    return out


def roofline_for(pkg, d, peak, peak_src, kernel_name, workload, world, variant=0):
    it_b, k1_b = d.bytes()
    if variant == 0 and d.one_kernel_iteration():
        # dense path: the whole classic iteration is ONE launch (k_cg_dense_iter) -- its algorithmic
        # bytes are the contract bytes of an iteration (A + 88 n)
        kernel_name, alg, which = "k_cg_dense_iter", it_b, 6
        what = "the one-kernel iteration (matvec, both reductions, both vector updates)"
    else:
        alg, which, what = k1_b, 1, "the matvec kernel"
    t1 = d.time_kernel(which, 30)
    ach = alg / (t1 * 1e-3) / 1e9
    return {
        "bound": "hbm", "kernel": kernel_name, "achieved": round(ach, 1), "peak": peak, "unit": "GB/s",
        "frac": round(ach / peak, 4), "traffic": traffic_from_profiles(kernel_name, workload, world),
        "algorithmic_bytes_per_launch": alg, "avg_launch_ms": round(t1, 5),
        "peak_source": peak_src,
        "how": "CUDA events on the library's solve stream around 30 back-to-back launches of "
               + what + " (this rank's block), after 3 warm-ups",
    }


# -------------------------------------------------------------------------------------------------
# end-to-end through the reference-facing C ABI with host buffers
# -------------------------------------------------------------------------------------------------


def operator_api_loop(pkg, A, b, bnp, bs, J):
    """The reference's own loop (cg.c:104-135) kept by the caller and run operator by operator
    against libcgb200.so in resident mode -- what a user gets who relinks mv_ops but keeps their
    cg.c.  Per iteration, as the reference issues them: mv_mult, 4 dot_product (2 answered from the
    stored v.v of an unchanged vector), 3 sv_mult, 2 vec_add, vec_sub, 2 mv_deep_copy; b in and x out
    every solve.  `value` is wall clock over everything; `loop_only` leaves out the per-solve set-up
    the reference's structure imposes on the host (deep copies of b on the host, cg.c:107-108, and
    the uploads they cause) and the final download."""
    L = pkg.lib()
    assert L.cgb200_set_mode(pkg.RESIDENT) == 0
    iters = 0
    loop_s = 0.0
    try:
        t = time.perf_counter()
        for bi in bs:
            bnp[:] = bi
            L.cgb200_invalidate(C.byref(b))
            x_new = L.new_mv_struct_with_size(b.size)
            r = L.mv_deep_copy(C.byref(b))
            p = L.mv_deep_copy(r)
            sv, sres = pkg.MvSparseP(), pkg.MvSparseP()
            k = 0
            while True:
                if k == 1:                     # everything is on the device from here on
                    tl = time.perf_counter()
                assert L.mv_mult(C.byref(A), p, C.byref(sv)) == 0, pkg.last_error()
                alpha = L.dot_product(r, r) / L.dot_product(p, sv)
                L.sv_mult(alpha, p, C.byref(sres))
                x_old = L.mv_deep_copy(x_new)
                L.vec_add(x_old, sres, C.byref(x_new))
                r_prev = L.mv_deep_copy(r)
                L.sv_mult(alpha, sv, C.byref(sres))
                L.vec_sub(r, sres, C.byref(r))
                iters += 1
                if k == J - 1:
                    L.cgb200_free_all(x_old)
                    L.cgb200_free_all(r_prev)
                    break
                beta = L.dot_product(r, r) / L.dot_product(r_prev, r_prev)
                L.sv_mult(beta, p, C.byref(sres))
                L.vec_add(r, sres, C.byref(p))
                k += 1
                L.cgb200_free_all(x_old)       # the reference leaks these two per iteration
                L.cgb200_free_all(r_prev)
            rr_end = L.dot_product(r, r)       # a read-back: the device has finished the loop
            loop_s += time.perf_counter() - tl
            assert L.cgb200_to_host(x_new) == 0, pkg.last_error()
            chk = float(x_new.contents.values[0])
            for o in (x_new, r, p, sv, sres):
                L.cgb200_free_all(o)
        sec = time.perf_counter() - t
    finally:
        L.cgb200_set_mode(pkg.EAGER)
    assert chk == chk and rr_end == rr_end
    return {"value": round(iters / sec, 2), "unit": UNIT, "solves": len(bs), "iterations_per_solve": J,
            "loop_only": {"value": round((iters - len(bs)) / loop_s, 2), "unit": UNIT,
                          "note": "iterations 2..J of every solve (vectors already on the device)"},
            "mode": "CGB200_RESIDENT",
            "note": "the reference's loop statement by statement through mv_mult / dot_product / sv_mult / vec_add / "
                    "vec_sub / mv_deep_copy; wall clock; b in and x out every solve"}


def e2e_host_structs(pkg, d, info, J, solves=8):
    """N = 1: cgb200_solve(A, b, &x) on struct __mv_sparse host arrays (pinned), every solve copying
    its right-hand side host->device and its solution device->host.  A is uploaded by the first
    call (reported as cold) and stays resident while its struct is unchanged."""
    import numpy as np
    import torch

    L = pkg.lib()
    n, nnz = d.rows, d.nnz
    vals = torch.empty(nnz, dtype=torch.float64, pin_memory=True)
    cols = torch.empty(nnz, dtype=torch.int32)
    rp = torch.empty(n + 1, dtype=torch.int32)
    bh = torch.empty(n, dtype=torch.float64, pin_memory=True)
    vnp, cnp, rnp, bnp = vals.numpy(), cols.numpy(), rp.numpy(), bh.numpy()
    assert L.cgb200_system_download(d.h, vnp.ctypes.data_as(pkg.c_double_p), cnp.ctypes.data_as(pkg.c_int_p),
                                    rnp.ctypes.data_as(pkg.c_int_p)) == 0, pkg.last_error()
    assert L.cgb200_system_get_b(d.h, bnp.ctypes.data_as(pkg.c_double_p)) == 0
    A = pkg.MvSparse(n, nnz, vnp.ctypes.data_as(pkg.c_double_p), cnp.ctypes.data_as(pkg.c_int_p),
                     rnp.ctypes.data_as(pkg.c_int_p))
    b = pkg.MvSparse(n, n, bnp.ctypes.data_as(pkg.c_double_p), None, None)
    o = pkg.Opts(200, 1e-10, 0, 0, 0) if info["kind"] == "dense" else pkg.Opts(J - 1, 0.0, 0, 0, 0)
    res = pkg.Result()

    def one():
        xp = pkg.MvSparseP()
        rc = L.cgb200_solve(C.byref(A), C.byref(b), C.byref(xp), C.byref(o), C.byref(res))
        assert rc == 0, pkg.last_error()
        chk = float(xp.contents.values[0])
        L.cgb200_free_all(xp)
        return res.iterations, chk

    t = time.perf_counter()
    it_cold, _ = one()                       # uploads A (8 B/nnz on the dense path) + b
    cold_s = time.perf_counter() - t
    rng = np.random.default_rng(3)
    bs = [rng.random(n) - 0.5 for _ in range(solves)]
    iters = 0
    t = time.perf_counter()
    for i in range(solves):
        bnp[:] = bs[i]                       # caller refreshes b in place; A untouched
        it, _ = one()
        iters += it
    warm_s = time.perf_counter() - t
    try:
        ops = operator_api_loop(pkg, A, b, bnp, bs, int(round(iters / solves)))
    except Exception as e:  # noqa: BLE001
        ops = {"error": repr(e)}
    L.cgb200_invalidate(C.byref(A))
    a_bytes = 8 * nnz if info["kind"] == "dense" else 12 * nnz + 4 * (n + 1)
    per_solve_iters = iters / solves
    return {
        "value": round(iters / warm_s, 2), "unit": UNIT,
        "h2d_bytes_per_step": round(8 * n / per_solve_iters, 1),
        "d2h_bytes_per_step": round(8 * n / per_solve_iters, 1),
        "api": "cgb200_solve(struct __mv_sparse* A, b, &x) == conj_grad with tol 1e-10; host arrays pinned",
        "solves": solves, "iterations_per_solve": per_solve_iters,
        "per_solve": {"h2d_bytes": 8 * n, "d2h_bytes": 8 * n, "ms": round(warm_s / solves * 1e3, 3)},
        "cold_first_call": {"value": round(it_cold / cold_s, 3), "unit": UNIT, "seconds": round(cold_s, 3),
                            "h2d_bytes": a_bytes + 8 * n, "note": "includes uploading A"},
        "note": "A stays device-resident between calls while its struct is unchanged (fingerprinted); every solve copies b in and x out",
        "operator_api": ops,
    }


def e2e_row_blocks(pkg, name, rank, world, J, solves=8):
    """N > 1: each rank hands its row block as host CSR arrays to cgb200_system_from_csr, then per
    solve copies its slice of b in and its slice of x out."""
    import numpy as np
    import torch

    L = pkg.lib()
    d, info = make_system(pkg, name, rank, world)
    n, nnz = d.rows, d.nnz
    vals = torch.empty(nnz, dtype=torch.float64, pin_memory=True).numpy()
    cols = np.empty(nnz, dtype=np.int32)
    rp = np.empty(n + 1, dtype=np.int32)
    bh = torch.empty(n, dtype=torch.float64, pin_memory=True).numpy()
    xh = torch.empty(n, dtype=torch.float64, pin_memory=True).numpy()
    assert L.cgb200_system_download(d.h, vals.ctypes.data_as(pkg.c_double_p), cols.ctypes.data_as(pkg.c_int_p),
                                    rp.ctypes.data_as(pkg.c_int_p)) == 0, pkg.last_error()
    assert L.cgb200_system_get_b(d.h, bh.ctypes.data_as(pkg.c_double_p)) == 0
    row0 = rank * n
    d.free()
    dist_sync(world)
    t = time.perf_counter()
    h = pkg.DeviceSystem.from_csr(info["n"], row0, rp, cols, vals, bh)
    dist_sync(world)
    cold_s = time.perf_counter() - t
    tol, mi = (1e-10, 200) if info["kind"] == "dense" else (0.0, J - 1)
    h.solve(mi, tol=tol)
    rng = np.random.default_rng(3 + rank)
    bs = [rng.random(n) - 0.5 for _ in range(solves)]
    dist_sync(world)
    iters = 0
    t = time.perf_counter()
    for i in range(solves):
        bh[:] = bs[i]
        h.set_rhs(bh)
        r = h.solve(mi, tol=tol)
        assert L.cgb200_system_get_x(h.h, xh.ctypes.data_as(pkg.c_double_p)) == 0
        iters += r.iterations
    dist_sync(world)
    warm_s = max_over_ranks(time.perf_counter() - t, world)
    h.free()
    per_solve_iters = iters / solves
    return {
        "value": round(iters / warm_s, 2), "unit": UNIT,
        "h2d_bytes_per_step": round(8 * n * world / per_solve_iters, 1),
        "d2h_bytes_per_step": round(8 * n * world / per_solve_iters, 1),
        "api": "cgb200_system_from_csr(host row block) + set_rhs / solve / get_x per solve, all ranks; host arrays pinned",
        "solves": solves, "iterations_per_solve": per_solve_iters,
        "cold_upload_seconds": round(cold_s, 3),
    }


# -------------------------------------------------------------------------------------------------
# CPU legs (the only places that execute oracle/)
# -------------------------------------------------------------------------------------------------


def cpu_reference_run(workload, budget_s, want_steps, host_arrays=None):
    """Times the reference's own CPU implementation on the host cores.

    oracle/_ref/libcgref.so (the UNMODIFIED reference, gcc -g as its Makefile builds it) when
    present, else the oracle port.  The reference is single-threaded, cores = 1.
    Returns (cpu_baseline dict, steps_run, seconds)."""
    import numpy as np

    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import helpers as H

    assert workload == "dense32768", "the reference arm is defined on the headline workload"
    n = 32768
    if host_arrays is None:
        log("[cpu] generating dense n=32768 CSR on the host (oracle generator)...")
        s = H.dense(n)
        rp, ci, va, b = s.row_ptr, s.cols, s.vals, s.b
    else:
        rp, ci, va, b = host_arrays
    have_ref = H.have_ref()
    kind = "reference" if have_ref else "port"

    def run(max_iter):
        t = time.perf_counter()
        if have_ref:
            R = H.ref()
            A = H.MvSparse(n, len(va), H.dptr(va), H.iptr(ci), H.iptr(rp))
            bv = H.MvSparse(n, n, H.dptr(b), None, None)
            xp = H.MvSparseP()
            assert R.conj_grad(max_iter, C.byref(A), C.byref(bv), C.byref(xp)) == 0
            x0 = float(xp.contents.values[0])
        else:
            x = np.empty(n)
            H.oracle().oracle_conj_grad(n, H.iptr(rp), H.iptr(ci), H.dptr(va), H.dptr(b), max_iter, 0.0, 1,
                                        H.dptr(x), None, None)
            x0 = float(x[0])
        return time.perf_counter() - t, x0

    t1, _ = run(0)  # one iteration: calibration and warm-up
    m = int(max(1, min(want_steps, budget_s // max(t1, 1e-3))))
    m = min(m, 5)   # fixed-count CG on this matrix is past convergence after ~6 iterations
    tm, _ = run(m - 1)
    val = m / tm
    # beside it, for context: the same loop as a tuned CPU program would run it -- O(nnz) CSR matvec,
    # fused passes, OpenMP over all host cores (oracle_cg_iterations_mt; timing only, its reduction
    # order is not the reference's).  Reported, never used as the baseline of a ratio.
    port = None
    try:
        O = H.oracle()
        O.oracle_cg_iterations_mt.restype = C.c_double
        O.oracle_cg_iterations_mt.argtypes = [C.c_int, H.c_int_p, H.c_int_p, H.c_double_p, H.c_double_p, C.c_int,
                                              H.c_double_p]
        xs = np.empty(n)
        O.oracle_cg_iterations_mt(n, H.iptr(rp), H.iptr(ci), H.dptr(va), H.dptr(b), 1, H.dptr(xs))  # warm-up
        tp = time.perf_counter()
        O.oracle_cg_iterations_mt(n, H.iptr(rp), H.iptr(ci), H.dptr(va), H.dptr(b), 5, H.dptr(xs))
        tp = time.perf_counter() - tp
        port = {"value": round(5 / tp, 3), "unit": UNIT, "cores": int(O.oracle_num_threads()), "kind": "port",
                "sample": "5 iterations of the O(nnz) CSR restatement with fused passes, OpenMP over all host "
                          "threads (oracle/cg_oracle.c: oracle_cg_iterations_mt, -O2), same system",
                "seconds": round(tp, 3)}
    except Exception as e:  # noqa: BLE001
        port = {"error": repr(e)}
    cb = {
        "value": round(val, 5), "unit": UNIT, "cores": 1, "kind": kind, "tuned_port_all_cores": port,
        "sample": f"{m} iteration(s) (max_iter={m - 1}) of {'the unmodified reference conj_grad (oracle/_ref, gcc -g as its Makefile)' if have_ref else 'the oracle port (oracle/cg_oracle.c, -O2)'} on the full dense n=32768 CSR system, after a 1-iteration warm-up ({t1:.1f} s)",
        "seconds": round(tm, 2),
    }
    return cb, m, tm


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    cb, m, tm = cpu_reference_run(args.workload, budget_s=100.0, want_steps=args.steps)
    line = {
        "impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": m, "warmup": 1, "ms_per_step": round(tm / m * 1e3, 2), "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args.workload, 1),
        "cpu_baseline": cb,
        "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "reference's own CPU implementation on the box's host cores; steps bounded to keep the run within minutes",
    }
    print(json.dumps(line), flush=True)
    return 0


def workload_config(name, world):
    cfgs = {
        "dense32768": {"workload": "dense SPD n=32768 fp64 (diagonally dominant, synthetic), CG to rel. residual 1e-10 (BASELINE configs[1])",
                       "n": 32768, "nnz": 32768 ** 2, "matvec": "values-only row-block GEMV", "partition": f"row blocks x{world}",
                       "l2": "inputs larger than L2 (8.6 GB matrix streamed every iteration)"},
        "lap2d4096": {"workload": "2D 5-point Poisson 4096x4096 CSR fp64 (BASELINE configs[2])", "n": 4096 ** 2,
                      "matvec": "CSR stream SpMV", "l2": "inputs larger than L2 (2.5 GB per iteration)"},
        "lap3d512": {"workload": "3D 7-point Laplacian 512^3 CSR fp64 (BASELINE configs[3])", "n": 512 ** 3,
                     "matvec": "CSR stream SpMV", "partition": f"z-slabs x{world}, halo exchange", "l2": "inputs larger than L2"},
        "lap3d384w": {"workload": "3D 7-point Laplacian 384^3 per GPU, weak scaling (BASELINE configs[4])",
                      "n": 384 ** 3 * world, "matvec": "CSR stream SpMV", "partition": f"z-slabs x{world}, halo exchange",
                      "l2": "inputs larger than L2"},
    }
    return cfgs[name]


# -------------------------------------------------------------------------------------------------


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--workload", default="dense32768")
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--variant", default="classic", choices=["classic", "single-reduction"],
                    help="recurrence of the headline run (the reference's classic one by default)")
    ap.add_argument("--no-extras", action="store_true", help="skip the other BASELINE configs")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-tts", action="store_true", help="skip the time-to-solution solves (profiling runs)")
    ap.add_argument("--cpu-leg", action="store_true", help="internal: print the cpu_baseline object only")
    args = ap.parse_args()
    if args.impl == "reference":
        return reference_arm(args)
    if args.cpu_leg:
        cb, _, _ = cpu_reference_run(args.workload, budget_s=20.0, want_steps=2)
        print(json.dumps(cb), flush=True)
        return 0

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("--gpus N > 1 must be launched with torch.distributed.run --nproc-per-node N")
        args.gpus = world

    import torch

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product has no CPU fallback)")
    torch.cuda.set_device(local_rank)
    import __graft_entry__ as g

    pkg = g.load_package()
    L = pkg.lib()
    assert L.cgb200_init(local_rank) == 0, pkg.last_error()
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        idbuf = (C.c_ubyte * 128)()
        if rank == 0:
            assert L.cgb200_dist_unique_id(idbuf) == 0, pkg.last_error()
        t = torch.tensor(list(bytes(idbuf)), dtype=torch.uint8, device="cuda")
        dist.broadcast(t, 0)
        raw = bytes(t.cpu().tolist())
        idbuf = (C.c_ubyte * 128).from_buffer_copy(raw)
        assert L.cgb200_dist_init(rank, world, idbuf) == 0, pkg.last_error()

    peak, peak_src = measured_peak()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    time.sleep(0.3)

    # ---- headline ----
    variant = 1 if args.variant == "single-reduction" else 0
    d, info = make_system(pkg, args.workload, rank, world)
    comm = "single" if world == 1 else ("p2p" if L.cgb200_system_comm_is_p2p(d.h) else "nccl")
    runner = Runner(pkg, d, info, variant)
    tm = timed_steps(pkg, runner, args.steps, args.warmup, world)
    clocks = sampler.summary(tm["t0"], tm["t1"]) if sampler else None
    it_b, k1_b = d.bytes()
    value = tm["steps"] / (tm["ms"] * 1e-3)
    kname = "k_gemv_units" if info["kind"] == "dense" else "k_spmv_stream"
    roof = roofline_for(pkg, d, peak, peak_src, kname, args.workload, world, variant)
    iter_gbs = it_b / (tm["ms"] / tm["steps"] * 1e-3) / 1e9
    true_rr = d.true_relres() if info["kind"] != "dense" else None
    line = {
        "metric": METRIC, "value": round(value, 2), "unit": UNIT, "n_gpus": world, "steps": tm["steps"],
        "warmup": tm["warmup"], "ms_per_step": round(tm["ms"] / tm["steps"], 5), "higher_is_better": True,
        "scaling": info["scaling"], "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": dict(workload_config(args.workload, world), iterations_per_solve=runner.J,
                       variant=args.variant, comm=comm,
                       kernels_per_iteration=1 if roof["kernel"] == "k_cg_dense_iter" else (3 if variant == 0 else 2)),
        "roofline": roof,
        "iteration": {"contract_bytes_per_rank": it_b, "achieved_gbs_per_gpu": round(iter_gbs, 1),
                      "frac_of_peak": round(iter_gbs / peak, 4),
                      "note": "contract bytes/iteration (SURVEY 8d: A + 88 n) / device time per iteration, per GPU"},
        "gpu_launches": tm["launches"],
        "clocks": clocks,
    }
    J = runner.J
    try:
        if not args.no_tts:
            line["time_to_solution"] = time_to_solution(pkg, d, info, world, variant,
                                                        fixed=500 if args.workload == "lap3d384w" else None)
    except Exception as e:  # noqa: BLE001
        line["time_to_solution"] = {"error": repr(e)}
    # the other recurrence on the same system, same K (explains, does not replace, the headline)
    try:
        other = Runner(pkg, d, info, 1 - variant)
        to = timed_steps(pkg, other, args.steps, args.warmup, world)
        line["other_variant"] = {"variant": "classic" if variant else "single-reduction",
                                 "value": round(to["steps"] / (to["ms"] * 1e-3), 2), "unit": UNIT,
                                 "ms_per_step": round(to["ms"] / to["steps"], 5), "gpu_launches": to["launches"]}
    except Exception as e:
        line["other_variant"] = {"error": repr(e)}
    d.free()

    # ---- end to end with host buffers ----
    if not args.no_e2e:
        try:
            if world == 1:
                d2, info2 = make_system(pkg, args.workload, 0, 1)
                J2 = J if J else 50
                line["e2e"] = e2e_host_structs(pkg, d2, info2, J2)
                host_arrays = None
                d2.free()
            else:
                line["e2e"] = e2e_row_blocks(pkg, args.workload, rank, world, J if J else 50)
        except Exception as e:  # an honest gap beats a crashed bench line
            line["e2e"] = {"value": None, "unit": UNIT, "error": repr(e)}

    # ---- the other BASELINE configs, same run ----
    if not args.no_extras:
        extras = {}
        names = ["lap2d4096", "lap3d512", "lap3d384w"] if world == 1 else ["lap3d512", "lap3d384w"]
        for nm in names:
            if nm == args.workload:
                continue
            try:
                dx, ix = make_system(pkg, nm, rank, world)
                rx = Runner(pkg, dx, ix, variant)
                k = 300 if nm == "lap2d4096" else 100
                tx = timed_steps(pkg, rx, k, 5, world)
                ib, _ = dx.bytes()
                gbs = ib / (tx["ms"] / tx["steps"] * 1e-3) / 1e9
                ex = {"value": round(tx["steps"] / (tx["ms"] * 1e-3), 2), "unit": UNIT, "steps": tx["steps"],
                      "ms_per_step": round(tx["ms"] / tx["steps"], 5), "scaling": ix["scaling"],
                      "achieved_gbs_per_gpu": round(gbs, 1), "frac_of_peak": round(gbs / peak, 4),
                      "config": workload_config(nm, world)}
                ex["roofline"] = roofline_for(pkg, dx, peak, peak_src, "k_spmv_stream", nm, world)
                try:
                    if not args.no_tts:
                        ex["time_to_solution"] = time_to_solution(pkg, dx, ix, world, variant,
                                                                  fixed=500 if nm == "lap3d384w" else None)
                except Exception as e:  # noqa: BLE001
                    ex["time_to_solution"] = {"error": repr(e)}
                ry = Runner(pkg, dx, ix, 1 - variant)
                ty = timed_steps(pkg, ry, k, 5, world)
                ex["other_variant"] = {"variant": "classic" if variant else "single-reduction",
                                       "value": round(ty["steps"] / (ty["ms"] * 1e-3), 2),
                                       "ms_per_step": round(ty["ms"] / ty["steps"], 5)}
                # the same system in the row-pattern dictionary format: reported beside,
                # never instead of, the contract figures above -- its iteration streams fewer bytes
                try:
                    csr_bytes = ib
                    if dx.use_pattern_format(True):
                        rp = Runner(pkg, dx, ix, variant)
                        tp = timed_steps(pkg, rp, k, 5, world)
                        pb, pk1 = dx.bytes()
                        ms_p = tp["ms"] / tp["steps"]
                        k1p = dx.time_kernel(1, 30)
                        ex["pattern_format"] = {
                            "value": round(tp["steps"] / (tp["ms"] * 1e-3), 2), "unit": UNIT,
                            "ms_per_step": round(ms_p, 5),
                            "speedup_vs_csr": round((tx["ms"] / tx["steps"]) / ms_p, 3),
                            "bytes_per_iteration": pb, "csr_contract_bytes_per_iteration": csr_bytes,
                            "achieved_gbs_per_gpu": round(pb / (ms_p * 1e-3) / 1e9, 1),
                            "frac_of_peak": round(pb / (ms_p * 1e-3) / 1e9 / peak, 4),
                            "matvec": {"kernel": "k_spmv_pattern", "algorithmic_bytes_per_launch": pk1,
                                       "avg_launch_ms": round(k1p, 5),
                                       "achieved": round(pk1 / (k1p * 1e-3) / 1e9, 1), "unit": "GB/s"},
                            "note": "one byte per row + shared-memory dictionary instead of 12 bytes per entry; same "
                                    "arithmetic, s = A p bit-identical to the CSR kernel (tests/test_gpu_parity.py)",
                        }
                        dx.use_pattern_format(False)
                    else:
                        ex["pattern_format"] = {"applicable": False, "why": pkg.last_error()}
                except Exception as e:  # noqa: BLE001
                    ex["pattern_format"] = {"error": repr(e)}
                extras[nm] = ex
                dx.free()
            except Exception as e:
                extras[nm] = {"error": repr(e)}
        if world == 1:
            try:
                extras["banded52269"] = long_row_leg(pkg, peak, peak_src, variant)
            except Exception as e:  # noqa: BLE001
                extras["banded52269"] = {"error": repr(e)}
        line["workloads"] = extras

    if sampler:
        sampler.stop()

    # ---- CPU baseline (rank 0, N = 1 only; the one leg that runs oracle/) ----
    if world == 1 and not args.no_cpu and args.workload == "dense32768":
        try:
            # a fresh process that never loads libcgb200.so: the reference must run its own operators
            out = subprocess.run([sys.executable, os.path.abspath(__file__), "--cpu-leg", "--workload", args.workload],
                                 stdout=subprocess.PIPE, text=True, timeout=900)
            line["cpu_baseline"] = json.loads(out.stdout.strip().splitlines()[-1])
        except Exception as e:
            line["cpu_baseline"] = {"value": None, "unit": UNIT, "error": repr(e)}

    if world > 1:
        import torch.distributed as dist

        L.cgb200_dist_finalize()
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        print(json.dumps(line), flush=True)
    return 0


if __name__ == "__main__":
    sys.exit(main())
