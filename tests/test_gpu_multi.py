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
    assert abs(two["relres"] - (rr / bb) ** 0.5) <= 1e-9 * (rr / bb) ** 0.5


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
        assert abs(j["relres"] - (rr / bb) ** 0.5) <= 1e-9 * (rr / bb) ** 0.5
    assert two["gpus"] == 2

