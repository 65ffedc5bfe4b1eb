"""Multi-GPU plumbing: one process per GPU (torch.distributed, NCCL on GPUs / gloo in CPU tests).

The task grid shards embarrassingly: tasks never talk to each other, the returns
matrices are replicated on every GPU, and there is exactly one exchange — the final gather of
result records.  `shard` is the static cyclic partition; `gather_records` the one collective
(one all-gather of packed per-task records, whatever the number of result tensors);
`gather_results` the same for a single tensor.
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


def shard_blocks(n_tasks: int, rank: int, world_size: int) -> np.ndarray:
    """Contiguous assignment: rank r takes the r-th of world_size nearly equal runs of consecutive tasks (the first
    n % world ranks one more).  For rolling windows of equal cost: a rank then needs only its own stretch of the history."""
    base, extra = divmod(n_tasks, world_size)
    lo = rank * base + min(rank, extra)
    return np.arange(lo, lo + base + (1 if rank < extra else 0), dtype=np.int64)


def _global_order(n_tasks: int, world_size: int, per: int, device, order: str = "cyclic") -> torch.Tensor:
    """Row of the rank-major gathered buffer [world][per] that holds global task i — cyclic: rank i % world, slot
    i // world; blocks: rank and slot of `shard_blocks`."""
    i = torch.arange(n_tasks, device=device)
    if order == "cyclic":
        return (i % world_size) * per + i // world_size
    if order != "blocks":
        raise ValueError(f"order {order!r}: 'cyclic' or 'blocks'")
    base, extra = divmod(n_tasks, world_size)
    split = extra * (base + 1)                                   # tasks below `split` sit on the ranks that hold base + 1
    r_hi = extra + (i - split) // max(base, 1)
    r = torch.where(i < split, i // (base + 1), r_hi)
    start = r * base + torch.clamp(r, max=extra)
    return r * per + (i - start)


def gather_results(local: torch.Tensor, n_tasks: int, rank: int | None = None, world_size: int | None = None,
                   order: str = "cyclic"):
    """All-gather per-task result rows (first dim = this rank's tasks in `shard` order — or `shard_blocks` order with
    order="blocks") and put them back in global task order.  Ragged shards are padded to ceil(n/world) rows."""
    r, w = world()
    rank = r if rank is None else rank
    world_size = w if world_size is None else world_size
    if world_size == 1:
        return local
    per = -(-n_tasks // world_size)
    pad = torch.zeros((per,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[:local.shape[0]] = local
    full = torch.empty((world_size * per,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(full, pad)
    return full.index_select(0, _global_order(n_tasks, world_size, per, local.device, order))


def pack_records(tensors) -> tuple[torch.Tensor, list]:
    """[n_local][...] tensors of any dtypes -> one uint8 matrix [n_local][record_bytes] (fields padded to 8 bytes) and
    the layout needed to take it apart again: [(offset, nbytes, dtype, trailing shape)], None entries kept as None."""
    cols, layout, off = [], [], 0
    n = None
    for t in tensors:
        if t is None:
            layout.append(None)
            continue
        n = t.shape[0] if n is None else n
        assert t.shape[0] == n, "every result tensor carries one row per local task"
        b = t.contiguous().reshape(n, -1).view(torch.uint8)           # [n][row bytes]
        nb = b.shape[1]
        padded = (nb + 7) & ~7
        if padded != nb:
            b = torch.cat([b, torch.zeros((n, padded - nb), dtype=torch.uint8, device=b.device)], dim=1)
        cols.append(b)
        layout.append((off, nb, t.dtype, tuple(t.shape[1:])))
        off += padded
    return torch.cat(cols, dim=1), layout


def unpack_records(rec: torch.Tensor, layout) -> list:
    out = []
    n = rec.shape[0]
    for item in layout:
        if item is None:
            out.append(None)
            continue
        off, nb, dtype, shape = item
        out.append(rec[:, off:off + nb].contiguous().view(dtype).reshape((n,) + shape))
    return out


def gather_records(tensors, n_tasks: int, rank: int | None = None, world_size: int | None = None) -> list:
    """The one collective of the multi-GPU path: every rank's per-task results (any number of tensors, first dim = the
    rank's tasks in `shard` order) travel as ONE all-gather of packed records and come back as tensors in global task
    order on every rank.  NCCL on GPUs (NVLink / NVSwitch), gloo on CPU tensors."""
    r, w = world()
    rank = r if rank is None else rank
    world_size = w if world_size is None else world_size
    if world_size == 1:
        return list(tensors)
    rec, layout = pack_records(tensors)
    per = -(-n_tasks // world_size)
    pad = torch.zeros((per, rec.shape[1]), dtype=torch.uint8, device=rec.device)
    pad[:rec.shape[0]] = rec
    full = torch.empty((world_size * per, rec.shape[1]), dtype=torch.uint8, device=rec.device)
    dist.all_gather_into_tensor(full, pad)
    return unpack_records(full.index_select(0, _global_order(n_tasks, world_size, per, rec.device)), layout)
