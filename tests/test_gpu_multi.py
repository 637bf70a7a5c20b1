"""Two ranks on two GPUs of one node over NCCL (skipped on a single-GPU box; run with `gpurun --gpus 2`): the task's
collectives (mask count all-reduce, seed broadcast with seed=None, spectra all-gather, group average) on each rank's own
device without the launcher having called torch.cuda.set_device."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_task_collectives_over_nccl_two_gpus():
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(ROOT, "tests", "_nccl_task_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "NCCL_TASK_OK 2" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]
