"""Replica data parallelism over independent utterances: one process per GPU, weights replicated,
utterance i -> rank i mod world, NO data-path collective.  `torch.distributed` is used only for the final host-side
gather of results and for the barrier / max-over-ranks timing of bench.py."""
from __future__ import annotations

import torch
import torch.distributed as dist


def shard_indices(n_items: int, rank: int, world: int) -> list[int]:
    """Round-robin assignment of utterance ids to replicas."""
    return list(range(rank, n_items, world))


def gather_ragged(local: dict[int, torch.Tensor], n_items: int, group=None) -> list[torch.Tensor] | None:
    """Gather per-utterance CPU tensors (ragged last dim) to rank 0, in utterance order.  Returns None elsewhere."""
    if not dist.is_available() or not dist.is_initialized():
        return [local[i] for i in range(n_items)]
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    payload = {i: t.cpu() for i, t in local.items()}
    gathered = [None] * world if rank == 0 else None
    dist.gather_object(payload, gathered, dst=0, group=group)
    if rank != 0:
        return None
    merged: dict[int, torch.Tensor] = {}
    for part in gathered:
        merged.update(part)
    missing = [i for i in range(n_items) if i not in merged]
    if missing:
        raise RuntimeError(f"utterances {missing} were not produced by any replica")
    return [merged[i] for i in range(n_items)]


def max_over_ranks(value: float, device: torch.device | str = "cpu", group=None) -> float:
    """Timing helper: device-timed duration, reduced with MAX over ranks."""
    if not dist.is_available() or not dist.is_initialized():
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())
