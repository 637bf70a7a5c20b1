"""Baseline sharding across ranks (one process per GPU) and the two collectives the path has.

Baselines are independent (the reference's loop, delay.py:1146-1178, has no communication),
so the data path needs NO collective: each rank owns a contiguous balanced block of the
flattened baseline axis, exactly like caput's ``MPIArray`` axis-0 distribution created at
delay.py:2298 / 794.  Collectives appear only after the chains:
  * ``gather_spectra``   all-gather of the per-baseline spectra when one rank must hold all;
  * ``group_average``    all-reduce(sum) of per-group partial sums + counts (BASELINE config 4).
Both go through ``torch.distributed`` (NCCL over NVLink on GPUs, gloo in the CPU tests).
"""

from __future__ import annotations

import numpy as np


def split_range(n, size, rank):
    """Contiguous balanced block of ``range(n)`` owned by ``rank`` (caput ``mpiutil.split_local``
    convention: the first ``n % size`` ranks get one extra element).  Returns (start, stop)."""
    base, rem = divmod(int(n), int(size))
    start = rank * base + min(rank, rem)
    return start, start + base + (1 if rank < rem else 0)


def _dist():
    import torch.distributed as dist

    return dist if dist.is_available() and dist.is_initialized() else None


def gather_spectra(local, nbase_total):
    """All-gather ``local`` ([B_local, N] torch tensor) into [nbase_total, N] on every rank."""
    import torch

    dist = _dist()
    if dist is None or dist.get_world_size() == 1:
        return local
    size = dist.get_world_size()
    counts = [split_range(nbase_total, size, r) for r in range(size)]
    bmax = max(b - a for a, b in counts)
    pad = torch.zeros((bmax, local.shape[1]), dtype=local.dtype, device=local.device)
    pad[: local.shape[0]] = local
    out = torch.empty((size * bmax, local.shape[1]), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, pad)
    return torch.cat([out[r * bmax : r * bmax + (b - a)] for r, (a, b) in enumerate(counts)], dim=0)


def group_average(local, global_ids, ngroup, group_size, weights=None):
    """Mean spectrum of each baseline group, identical on every rank.

    ``local`` [B_local, N]; ``global_ids`` [B_local] global baseline indices; group of a
    baseline = ``global_id // group_size``.  Baselines with ``weights == 0`` (e.g. failed
    chains) are left out.  One all-reduce of ``[ngroup, N+1]`` doubles.
    """
    import torch

    gid = torch.as_tensor(np.asarray(global_ids) // group_size, device=local.device, dtype=torch.long)
    w = torch.ones(local.shape[0], dtype=local.dtype, device=local.device) if weights is None else weights.to(local.dtype)
    acc = torch.zeros((ngroup, local.shape[1] + 1), dtype=local.dtype, device=local.device)
    acc[:, :-1].index_add_(0, gid, local * w[:, None])
    acc[:, -1].index_add_(0, gid, w)
    dist = _dist()
    if dist is not None and dist.get_world_size() > 1:
        dist.all_reduce(acc, op=dist.ReduceOp.SUM)
    cnt = acc[:, -1:]
    return torch.where(cnt > 0, acc[:, :-1] / cnt.clamp(min=1e-300), torch.zeros_like(acc[:, :-1])), cnt[:, 0]


def _collective_device(device=None):
    """Where a small collective's tensor lives: the CPU under gloo; under NCCL the task's own GPU (``device``), never
    the process-wide current device -- several ranks of one node would otherwise all land on cuda:0."""
    import torch

    dist = _dist()
    if dist is None or dist.get_backend() != "nccl":
        return torch.device("cpu")
    if device is None:
        device = torch.cuda.current_device()
    return torch.device("cuda", int(device))


def allreduce_count(n, device=None):
    """Scalar all-reduce used for the converged-baseline debug count (delay.py:1182)."""
    import torch

    dist = _dist()
    if dist is None or dist.get_world_size() == 1:
        return int(n)
    t = torch.tensor([int(n)], dtype=torch.int64, device=_collective_device(device))
    dist.all_reduce(t)
    return int(t.item())


def broadcast_int(value, src=0, device=None):
    """Make one Python int (< 2**63) identical on every rank (rank ``src`` wins)."""
    import torch

    dist = _dist()
    if dist is None or dist.get_world_size() == 1:
        return int(value)
    t = torch.tensor([int(value)], dtype=torch.int64, device=_collective_device(device))
    dist.broadcast(t, src=src)
    return int(t.item())
