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
