"""Instance sharding across the GPUs of one box.

Robots are independent, so the batch is cut into contiguous blocks of instance indices, one block per
rank / GPU, and the hot path needs NO collective.  The only cross-rank steps are host-side plumbing:
gathering the per-rank outputs into one array and reducing timings with a max over ranks.  They run
over ``torch.distributed`` (``nccl`` on the GPU box, ``gloo`` in the CPU tests).
"""
from __future__ import annotations

from typing import Optional, Tuple

import numpy as np

from .workloads import shard_range  # noqa: F401  (re-exported)


def gather_outputs(local: np.ndarray, n_global: int, rank: int, world: int, device: Optional[str] = None) -> Optional[np.ndarray]:
    """Gather per-rank ``[n_local, C]`` blocks into the global ``[n_global, C]`` array on rank 0
    (``None`` on the other ranks).  Blocks may have different lengths (ragged split)."""
    import torch
    import torch.distributed as dist
    if world == 1:
        return local
    sizes = [shard_range(n_global, r, world) for r in range(world)]
    maxlen = max(b - a for a, b in sizes)
    pad = np.zeros((maxlen,) + local.shape[1:], dtype=local.dtype)
    pad[: local.shape[0]] = local
    t = torch.from_numpy(pad)
    if device:
        t = t.to(device)
    bufs = [torch.empty_like(t) for _ in range(world)] if rank == 0 else None
    dist.gather(t, bufs, dst=0)
    if rank != 0:
        return None
    out = np.empty((n_global,) + local.shape[1:], dtype=local.dtype)
    for r, (a, b) in enumerate(sizes):
        out[a:b] = bufs[r].cpu().numpy()[: b - a]
    return out


def max_over_ranks(values, world: int, device: Optional[str] = None):
    """Element-wise max of a small vector of floats over all ranks (timing reduction)."""
    import torch
    import torch.distributed as dist
    t = torch.tensor(list(values), dtype=torch.float64)
    if device:
        t = t.to(device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return [float(x) for x in t.cpu()]


def sum_over_ranks(values, world: int, device: Optional[str] = None):
    import torch
    import torch.distributed as dist
    t = torch.tensor(list(values), dtype=torch.float64)
    if device:
        t = t.to(device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return [float(x) for x in t.cpu()]
