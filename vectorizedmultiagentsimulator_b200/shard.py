"""Sharding of the environment batch over the GPUs of one box: independent replicas, one process per GPU,
NO collective on the data path.  The only exchange is the optional end-of-rollout gather of per-env episode statistics.
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple


def shard_range(n_envs_total: int, rank: int, world_size: int) -> Tuple[int, int]:
    """Envs [begin, end) of `rank` when `n_envs_total` environments are split as evenly as possible."""
    base, rem = divmod(n_envs_total, world_size)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def shard_seed(seed: int, rank: int) -> int:
    """Replica `rank` is a reference World of its own (seed + rank): its broad-phase flags are over its own batch."""
    return seed + rank


def gather_stats(stats: Dict[str, "torch.Tensor"], dst: int = 0) -> Optional[Dict[str, "torch.Tensor"]]:
    """Gathers per-env statistics ([n_local, ...] tensors, equal n_local on all ranks is NOT required) to rank `dst`;
    returns {name: tensor over all envs in rank order} there and None elsewhere.  Works with nccl and gloo."""
    import torch
    import torch.distributed as dist

    if not dist.is_initialized() or dist.get_world_size() == 1:
        return dict(stats)
    rank, world = dist.get_rank(), dist.get_world_size()
    out = {}
    for name in sorted(stats):
        t = stats[name].contiguous()
        n = torch.tensor([t.shape[0]], device=t.device, dtype=torch.int64)
        sizes = [torch.zeros_like(n) for _ in range(world)]
        dist.all_gather(sizes, n)
        cap = int(max(int(s.item()) for s in sizes))
        pad = torch.zeros((cap,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        pad[: t.shape[0]] = t
        bufs = [torch.zeros_like(pad) for _ in range(world)] if rank == dst else None
        dist.gather(pad, bufs, dst=dst)
        if rank == dst:
            out[name] = torch.cat([b[: int(s.item())] for b, s in zip(bufs, sizes)], 0)
    return out if rank == dst else None
