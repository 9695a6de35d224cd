"""Real multi-GPU tests (>= 2 B200s, NCCL bootstrap + our P2P kernels), launched with torchrun the
way the reference's ``gpu_parallel`` tests use ``elastic_launch``."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SCRIPT = os.path.join(ROOT, "tests", "fixtures", "dist_check.py")

pytestmark = [pytest.mark.gpu, pytest.mark.gpu_parallel]


def _run(which: str, nproc: int, timeout: int = 240) -> None:
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nproc}", "--master-addr", "127.0.0.1",
           "--master-port", "29617", SCRIPT, which]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=timeout)
    assert r.returncode == 0 and "DIST_CHECK_OK" in r.stdout, r.stdout[-4000:]


def _ngpu() -> int:
    import torch

    return torch.cuda.device_count()


@pytest.mark.parametrize("which", ["allreduce", "allreduce_nvls", "engine", "gan", "amp", "ddp_hook", "resnet", "gemm_comm", "gpt_tp", "zero3", "zero3_engine"])
def test_two_ranks(which):
    _run(which, 2)


def test_all_ranks_allreduce_and_engine():
    n = _ngpu()
    if n < 4:
        pytest.skip("needs >= 4 GPUs")
    _run("allreduce", n)
    _run("allreduce_nvls", n)
    _run("engine", n)
