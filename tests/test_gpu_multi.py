"""Multi-GPU parity on the box: tests/dist_check.py under torch.distributed.run, one process per
GPU, in both exchange modes (peer memory over NVLink, and NCCL between the kernels).  Needs at
least two visible GPUs; with one GPU only the single-rank pass runs (the N > 1 host logic is
covered on CPU by tests/test_partition_gloo.py)."""
import os
import socket
import subprocess
import sys

import pytest

import helpers as H

pytestmark = pytest.mark.gpu


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _ngpus():
    import torch

    return torch.cuda.device_count()


def _run(nproc, comm):
    env = dict(os.environ, CGB200_COMM=comm)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}",
           "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
           os.path.join(H.ROOT, "tests", "dist_check.py")]
    out = subprocess.run(cmd, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    return out.returncode, out.stdout


@pytest.mark.parametrize("comm", ["p2p", "nccl"])
def test_two_ranks_match_the_oracle(gpu_lib, comm):
    if _ngpus() < 2:
        pytest.skip("one GPU visible: multi-rank parity needs two")
    rc, log = _run(2, comm)
    assert rc == 0 and "FAIL" not in log, log[-4000:]
    assert f"comm {comm}" in log, log[-2000:]


def test_single_rank_pass_of_the_multi_rank_checker(gpu_lib):
    rc, log = _run(1, "p2p")
    assert rc == 0 and "FAIL" not in log, log[-4000:]


# ---------------------------------------------------------------------------------------------
# the cg driver's own rank launcher (driver/cg_main.c: fork, NCCL id through pipes, one process per
# GPU, every rank taking its row block of the system)
# ---------------------------------------------------------------------------------------------


def _cg(pkg, *args, **env):
    import json

    out = subprocess.run([pkg.CG_DRIVER, *args], env=dict(os.environ, **env), stdout=subprocess.PIPE,
                         stderr=subprocess.STDOUT, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-3000:]
    js = [ln for ln in out.stdout.splitlines() if ln.startswith("{")]
    return out.stdout, (json.loads(js[-1]) if js else None)


@pytest.mark.parametrize("comm", ["p2p", "nccl"])
def test_cg_driver_two_gpus_on_a_generated_3d_laplacian(pkg, gpu_lib, comm):
    if _ngpus() < 2:
        pytest.skip("one GPU visible")
    s = H.lap(48, 48, 48)
    _, one = _cg(pkg, "--gen", "lap3d:48", "400", "y", "--tol", "1e-10", "--json")
    _, two = _cg(pkg, "--gen", "lap3d:48", "400", "y", "--tol", "1e-10", "--json", "--gpus", "2", CGB200_COMM=comm)
    _, it, _, st = H.oracle_cg(s, 400, tol=1e-10, guard=1)
    assert one["gpus"] == 1 and two["gpus"] == 2 and st == 1
    for j in (one, two):
        assert j["status"] == pkg.CONVERGED and j["relres"] <= 1e-10 and abs(j["iterations"] - it) <= 2
    # fixed count, the reference's semantics: 31 updates on either side
    _, two = _cg(pkg, "--gen", "lap3d:48", "30", "y", "--json", "--gpus", "2", CGB200_COMM=comm)
    assert two["iterations"] == 31 and two["status"] == pkg.MAXITER
    rr = H.oracle_cg(s, 30)[2]
    bb = float(s.b @ s.b)
    assert abs(two["relres"] - (rr / bb) ** 0.5) <= 1e-6 * (rr / bb) ** 0.5   # (the JSON prints 7 digits)


def test_cg_driver_two_gpus_on_an_input_file(pkg, gpu_lib, tmp_path):
    # every rank reads the file and takes rows [r0, r0 + nl): a slice of the global row_ptr with the
    # global cols / vals arrays (include/cgb200.h: cgb200_system_from_csr)
    if _ngpus() < 2:
        pytest.skip("one GPU visible")
    s = H.banded(1501, 40)          # odd size: uneven row blocks
    p = tmp_path / "banded.txt"
    H.write_text_format(s, str(p))
    _, one = _cg(pkg, str(p), "25", "y", "--json")
    _, two = _cg(pkg, str(p), "25", "y", "--json", "--gpus", "2")
    rr = H.oracle_cg(s, 25)[2]
    bb = float(s.b @ s.b)
    for j in (one, two):
        assert j["iterations"] == 26 and j["status"] == pkg.MAXITER
        assert abs(j["relres"] - (rr / bb) ** 0.5) <= 1e-6 * (rr / bb) ** 0.5   # (the JSON prints 7 digits)
    assert two["gpus"] == 2


# ---------------------------------------------------------------------------------------------
# several GPUs behind the reference's own entry point: CGB200_GPUS=N (include/cgb200.h,
# csrc/team.cu).  Fresh interpreters: the variable is read once per process.
# ---------------------------------------------------------------------------------------------

TEAM_SCRIPT = r"""
import ctypes as C, json, os, sys
import numpy as np
sys.path.insert(0, os.path.join(%(root)r, "tests")); sys.path.insert(0, %(root)r)
import __graft_entry__ as g, helpers as H
pkg = g.load_package(); L = pkg.lib()
out = {}
def solve(s, M, tol=0.0, flags=0, x0=None):
    A = pkg.MvSparse(s.n, s.nnz, H.dptr(s.vals), H.iptr(s.cols), H.iptr(s.row_ptr))
    b = pkg.MvSparse(s.n, s.n, H.dptr(s.b), None, None)
    xp = pkg.MvSparseP()
    if x0 is not None:
        xv = pkg.MvSparse(s.n, s.n, H.dptr(x0), None, None); xp = C.pointer(xv)
    o = pkg.Opts(M, tol, 0, 0, flags); r = pkg.Result()
    assert L.cgb200_solve(C.byref(A), C.byref(b), C.byref(xp), C.byref(o), C.byref(r)) == 0, pkg.last_error()
    return np.ctypeslib.as_array(xp.contents.values, shape=(s.n,)).copy(), r, A
rel = lambda x, y: float(np.linalg.norm(x - y) / np.linalg.norm(y))
# 1) conj_grad itself, fixed count: 3D Laplacian (neighbour halos), uneven banded blocks, dense (all-gather)
for name, s, M in [("lap3d", H.lap(40, 36, 30), 60), ("banded", H.banded(3001, 70), 9), ("dense", H.dense(1536), 4)]:
    A = pkg.MvSparse(s.n, s.nnz, H.dptr(s.vals), H.iptr(s.cols), H.iptr(s.row_ptr))
    b = pkg.MvSparse(s.n, s.n, H.dptr(s.b), None, None)
    xp = pkg.MvSparseP()
    assert L.conj_grad(M, C.byref(A), C.byref(b), C.byref(xp)) == 0, pkg.last_error()
    x = np.ctypeslib.as_array(xp.contents.values, shape=(s.n,)).copy()
    out[name] = {"rel": rel(x, H.oracle_cg(s, M)[0]), "gpus": L.cgb200_gpus_in_use()}
    # an in-place edit between two calls is seen by the team as by one GPU (default cache policy)
    if name == "banded":
        s.vals[s.row_ptr[1:] - 1] *= 1.5
        assert L.conj_grad(M, C.byref(A), C.byref(b), C.byref(xp)) == 0, pkg.last_error()
        x = np.ctypeslib.as_array(xp.contents.values, shape=(s.n,)).copy()
        out["banded_edited"] = {"rel": rel(x, H.oracle_cg(s, M)[0])}
# 2) tolerance stop, world reported, Jacobi + initial guess through the team
s = H.lap(64, 64)
x, r, A = solve(s, 2000, tol=1e-10)
xo, it, _, _ = H.oracle_cg(s, 2000, tol=1e-10, guard=1)
out["tol"] = {"rel": rel(x, xo), "iters": int(r.iterations), "iters_oracle": int(it), "world": int(r.world), "status": int(r.status),
              "relres_true": float(H.true_relres(s, x))}
x2, r2, _ = solve(s, 2000, tol=1e-10, flags=pkg.OPT_JACOBI | pkg.OPT_X0, x0=x.copy())
out["warm"] = {"iters": int(r2.iterations), "world": int(r2.world)}
# 3) the operators still run (on this process's GPU) while the team exists
v = H.rhs(s.n, 5); vv = pkg.MvSparse(s.n, s.n, H.dptr(v), None, None); o_ = pkg.MvSparseP()
assert L.mv_mult(C.byref(A), C.byref(vv), C.byref(o_)) == 0, pkg.last_error()
want = np.empty(s.n); H.oracle().oracle_spmv(s.n, H.iptr(s.row_ptr), H.iptr(s.cols), H.dptr(s.vals), H.dptr(v), H.dptr(want))
out["mv_mult_bits"] = bool(np.array_equal(np.ctypeslib.as_array(o_.contents.values, shape=(s.n,)), want))
L.cgb200_shutdown()
print("TEAM_JSON " + json.dumps(out))
"""


def _team_run(ngpu, **env):
    import json

    code = TEAM_SCRIPT % {"root": H.ROOT}
    out = subprocess.run([sys.executable, "-c", code], env=dict(os.environ, CGB200_GPUS=str(ngpu), **env),
                         stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-4000:]
    line = [ln for ln in out.stdout.splitlines() if ln.startswith("TEAM_JSON ")]
    assert line, out.stdout[-2000:]
    return json.loads(line[-1][len("TEAM_JSON "):]), out.stdout


@pytest.mark.parametrize("comm", ["p2p", "nccl"])
def test_conj_grad_on_a_gpu_team_matches_the_oracle(gpu_lib, comm):
    if _ngpus() < 2:
        pytest.skip("one GPU visible")
    r, log = _team_run(2, CGB200_COMM=comm)
    for k in ("lap3d", "banded", "banded_edited", "dense"):
        assert r[k]["rel"] <= 1e-8, (k, r[k])
    assert r["lap3d"]["gpus"] == 2 and r["dense"]["gpus"] == 2
    assert r["tol"]["world"] == 2 and r["tol"]["status"] == 2 and abs(r["tol"]["iters"] - r["tol"]["iters_oracle"]) <= 2
    assert r["tol"]["rel"] <= 1e-8 and r["tol"]["relres_true"] <= 1.05e-10
    assert r["warm"]["world"] == 2 and r["warm"]["iters"] <= 2     # a converged guess stops at once
    assert r["mv_mult_bits"]


def test_gpu_team_request_beyond_the_visible_gpus_falls_back_to_one(gpu_lib):
    r, log = _team_run(_ngpus() + 1)
    assert "solving on one GPU" in log
    assert r["lap3d"]["gpus"] == 1 and r["lap3d"]["rel"] <= 1e-8 and r["tol"]["world"] == 1


def test_cg_driver_reference_flow_runs_conj_grad_on_two_gpus(pkg, gpu_lib):
    # the reference's UNMODIFIED cg.c (oracle/_ref/cg_on_b200) with CGB200_GPUS=2: its conj_grad... is its
    # own loop over the operators (single GPU by construction); the driver `cg`, whose solve IS conj_grad,
    # must reproduce the reference CLI's golden stdout on two GPUs
    if _ngpus() < 2:
        pytest.skip("one GPU visible")
    inp = os.path.join(H.GOLDEN, "config1_lap2d_32x32.txt")
    want = [ln for ln in open(os.path.join(H.GOLDEN, "config1_cg_stdout.txt")).read().splitlines()
            if not ln.startswith("\tValues:") and not ln.startswith("CG took")]
    out = subprocess.run([pkg.CG_DRIVER, inp, "30"], env=dict(os.environ, CGB200_GPUS="2", CGB200_VERBOSE="1"),
                         stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    assert out.returncode == 0, out.stdout[-2000:]
    got = [ln for ln in out.stdout.splitlines() if not ln.startswith("\tValues:") and not ln.startswith("CG took")
           and not ln.startswith("[cgb200]") and not ln.startswith("NCCL version")]   # (NCCL's own banner under NCCL_DEBUG)
    assert len(got) == len(want)
    for g_, w_ in zip(got, want):
        assert g_ == w_ or abs(float(g_) - float(w_)) <= 1.0001e-6, (g_, w_)

