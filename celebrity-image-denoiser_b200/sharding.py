"""Batch sharding across the GPUs of one box: one process per GPU, contiguous image ranges, no
collective on the inference path (each image is an independent unit)."""
from __future__ import annotations

from typing import List, Tuple

import torch


def shard_range(n_items: int, rank: int, world_size: int) -> Tuple[int, int]:
    """[start, stop) of the contiguous shard `rank` owns; the first n_items % world_size ranks
    get one extra item, so shards differ by at most one and concatenate back in rank order."""
    if world_size < 1 or not (0 <= rank < world_size):
        raise ValueError(f"bad rank/world_size {rank}/{world_size}")
    base, extra = divmod(n_items, world_size)
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


def shard_sizes(n_items: int, world_size: int) -> List[int]:
    return [shard_range(n_items, r, world_size)[1] - shard_range(n_items, r, world_size)[0] for r in range(world_size)]


def gather_shards(local: torch.Tensor, n_items: int, group=None) -> torch.Tensor:
    """Reassemble per-rank output shards (dim 0) on every rank — used by tests / offline jobs that
    want the whole result in one place; the serving path never calls it."""
    import torch.distributed as dist

    world = dist.get_world_size(group)
    sizes = shard_sizes(n_items, world)
    pad = max(sizes)
    buf = torch.zeros((pad, *local.shape[1:]), dtype=local.dtype, device=local.device)
    buf[: local.shape[0]] = local
    outs = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(outs, buf, group=group)
    return torch.cat([o[:s] for o, s in zip(outs, sizes)], dim=0)
