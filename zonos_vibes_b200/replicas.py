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


def generate_sharded(model, prefixes: list, max_new_tokens: int = 86 * 30, *, max_batch: int = 64, cfg_scale: float = 2.0,
                     sampling_params: dict | None = None, seed: int = 0, decode_audio: bool = True, group=None,
                     stats: dict | None = None) -> list | None:
    """BASELINE config 3: a batch of independent utterances over the replicas of one node.

    `prefixes[i]` is the `prefix_conditioning` of utterance i, `[2, Lc, d]` (row 0 conditional, row 1 unconditional,
    model.py:204-212).  Utterance i goes to rank `i mod world`; every rank runs ITS utterances through the batched decode
    path (`Zonos.generate(batch_size=...)`, at most `max_batch` at a time, utterances of equal Lc together) and the DAC
    decoder, copies the results to pinned host memory and the pieces are gathered on rank 0 in utterance order.  There
    is NO collective on the data path: replicas share nothing but the (replicated) weights.

    Returns on rank 0 the list of waveforms `[1, 512 * T_i]` fp32 (or of code tensors `[9, T_i]` when
    `decode_audio=False`), `None` on the other ranks.  `stats`, when given, receives `frames` (generated on this rank)
    and `decode_step_ms` (device-timed mean over this rank's chunks).
    """
    import torch.distributed as dist

    world = dist.get_world_size(group) if dist.is_available() and dist.is_initialized() else 1
    rank = dist.get_rank(group) if world > 1 else 0
    mine = shard_indices(len(prefixes), rank, world)
    by_lc: dict[int, list[int]] = {}
    for i in mine:
        by_lc.setdefault(int(prefixes[i].shape[1]), []).append(i)
    local: dict[int, torch.Tensor] = {}
    frames, step_ms = 0, []
    for _, ids in sorted(by_lc.items()):
        for c0 in range(0, len(ids), max_batch):
            chunk = ids[c0:c0 + max_batch]
            cond = torch.cat([prefixes[i][:1] for i in chunk])
            uncond = torch.cat([prefixes[i][1:] for i in chunk])
            codes = model.generate(torch.cat([cond, uncond]), max_new_tokens=max_new_tokens, cfg_scale=cfg_scale, batch_size=len(chunk),
                                   sampling_params=dict(min_p=0.1) if sampling_params is None else sampling_params,
                                   progress_bar=False, seed=seed + chunk[0])
            lens = list(model.last_lengths)
            if model.last_step_ms is not None:
                step_ms.append(model.last_step_ms)
            frames += sum(lens)
            out = model.autoencoder.decode(codes) if decode_audio else codes
            hop = 512 if decode_audio else 1
            host = torch.empty(out.shape, dtype=out.dtype).pin_memory() if out.is_cuda else out
            if out.is_cuda:
                host.copy_(out, non_blocking=True)
                torch.cuda.current_stream(out.device).synchronize()
            for j, i in enumerate(chunk):
                local[i] = host[j, ..., : hop * lens[j]].clone()
    if stats is not None:
        stats["frames"] = frames
        stats["decode_step_ms"] = sum(step_ms) / len(step_ms) if step_ms else None
    return gather_ragged(local, len(prefixes), group)
