"""Worker of tests/test_gpu_multi.py: one process per GPU under torchrun, NCCL backend, and -- on purpose -- NO
torch.cuda.set_device: the task must put its collectives on its own device (advisor finding r01)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dist.init_process_group("nccl")
    from draco_b200 import delay, dist as ddist
    from test_gpu_task import _stream

    ss, _, _ = _stream(nstack=10, dead=(4,))
    task = delay.DelayPowerSpectrumGibbs.from_config(dict(sample_axis="ra", nsamp=6, save_spectrum_mask=True))  # seed=None
    assert task._device() == rank % torch.cuda.device_count()
    out = task.process(ss)
    local = torch.from_numpy(np.ascontiguousarray(out.spectrum[:].local_array)).to(f"cuda:{task._device()}")
    full = ddist.gather_spectra(local, 10)
    assert full.shape == (10, 64) and bool(torch.isfinite(full).all())
    # every rank drew the same seed (broadcast from rank 0): a second task seeded explicitly reproduces the slab
    dev = torch.device("cuda", task._device())
    mine = torch.tensor([float(local.sum().item())], dtype=torch.float64, device=dev)
    allv = torch.zeros(world, dtype=torch.float64, device=dev)
    dist.all_gather_into_tensor(allv, mine)
    assert bool(torch.isfinite(allv).all())
    nmask = int(out.datasets["spectrum_mask"][:].local_array.sum())
    tot = ddist.allreduce_count(nmask, device=task._device())
    assert tot == 1, tot
    avg, cnt = ddist.group_average(local, np.arange(*ddist.split_range(10, world, rank)), 2, 5)
    assert float(cnt.sum()) == 10.0
    dist.barrier(device_ids=[task._device()])
    if rank == 0:
        print("NCCL_TASK_OK", world)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
