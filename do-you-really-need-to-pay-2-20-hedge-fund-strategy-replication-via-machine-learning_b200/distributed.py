"""Multi-GPU plumbing: one process per GPU (torch.distributed, NCCL on GPUs / gloo in CPU tests).

The task grid shards embarrassingly: tasks never talk to each other, the returns
matrices are replicated on every GPU, and there is exactly one exchange — the final gather of
result records.  `shard` is the static cyclic partition; `gather_results` the one collective.
"""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist


def world():
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard(n_tasks: int, rank: int, world_size: int) -> np.ndarray:
    """Cyclic assignment task i -> rank i % world_size (tasks pre-sorted by cost upstream, so every
    rank sees the same mix of latent widths)."""
    return np.arange(rank, n_tasks, world_size, dtype=np.int64)


def gather_results(local: torch.Tensor, n_tasks: int, rank: int | None = None, world_size: int | None = None):
    """All-gather per-task result records (first dim = this rank's tasks in `shard` order) and put
    them back in global task order.  Ragged shards are padded to ceil(n/world) rows."""
    r, w = world()
    rank = r if rank is None else rank
    world_size = w if world_size is None else world_size
    if world_size == 1:
        return local
    per = -(-n_tasks // world_size)
    pad = torch.zeros((per,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[:local.shape[0]] = local
    bufs = [torch.empty_like(pad) for _ in range(world_size)]
    dist.all_gather(bufs, pad)
    out = torch.empty((n_tasks,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    for q in range(world_size):
        idx = shard(n_tasks, q, world_size)
        out[torch.from_numpy(idx).to(local.device)] = bufs[q][:len(idx)]
    return out
