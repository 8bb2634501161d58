"""Instance sharding of a batch of independent controller instances over the GPUs of one box.

The path has no exchange step: instance ``b`` of the batch is evaluated entirely on one GPU with replicated keys,
tables and encoded diagonals.  The only collective is the gather of the result ciphertexts (north_star: "NCCL over
NVLink is used only for the final result gather").  One process per GPU; ``torch.distributed`` is plumbing.
"""
from __future__ import annotations


def shard_range(total: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous block [lo, hi) of instance indices owned by ``rank``.  Blocks differ by at most one instance and
    cover [0, total) in rank order, so a gather in rank order restores the batch order."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside a world of {world}")
    if total < 0:
        raise ValueError("negative instance count")
    base, extra = divmod(total, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_sizes(total: int, world: int) -> list[int]:
    return [hi - lo for lo, hi in (shard_range(total, r, world) for r in range(world))]


def gather_results(local, total: int, group=None, out=None, async_op: bool = False):
    """All-gather the per-rank result words (a tensor whose dim 0 is this rank's instances) into batch order.

    Equal shards use one ``all_gather_into_tensor`` (NCCL: a single NVLink/NVSwitch collective); ragged shards pad to
    the largest shard first.  Returns a tensor with ``total`` instances on every rank.

    ``out``: preallocated destination (equal shards only).  ``async_op=True`` (equal shards only) returns ``(out, work)``
    without making the calling stream wait: the gather of step k then overlaps the evaluation of step k+1, and the caller
    calls ``work.wait()`` before it reads ``out`` or reuses ``local``."""
    import torch
    import torch.distributed as dist

    world = dist.get_world_size(group)
    sizes = shard_sizes(total, world)
    if local.shape[0] != sizes[dist.get_rank(group)]:
        raise ValueError(f"rank holds {local.shape[0]} instances, expected {sizes[dist.get_rank(group)]}")
    if world == 1:
        return (local, None) if async_op else local
    big = max(sizes)
    if min(sizes) == big:
        if out is None:
            out = torch.empty((total,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
        work = dist.all_gather_into_tensor(out, local.contiguous(), group=group, async_op=async_op)
        return (out, work) if async_op else out
    if async_op or out is not None:
        raise ValueError("ragged shards: only the blocking gather into a fresh tensor is implemented")
    pad = torch.zeros((big,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[: local.shape[0]] = local
    out = torch.empty((world * big,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, pad, group=group)
    return torch.cat([out[r * big: r * big + sizes[r]] for r in range(world)])
