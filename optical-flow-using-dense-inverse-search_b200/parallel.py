"""Multi-GPU plumbing: one process per GPU, frame pairs sharded, no collective on the data path.

Frame pairs are independent (/root/reference/src/main.cpp:102 — the loop body shares no state), so a
job of P pairs on N ranks is N independent engines; `torch.distributed` (NCCL on the GPU box, gloo in
the CPU tests) carries only the barrier, the max-over-ranks of the device time and the gather of
run statistics.  Nothing here touches pixel data.
"""
from __future__ import annotations

from typing import Dict, List, Tuple


def shard_pairs(n_pairs: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous block [lo, hi) of pair indices for `rank`: sizes differ by at most one, earlier
    ranks get the longer blocks.  Contiguity matters for sequences — consecutive pairs share a frame,
    so a rank builds each pyramid once (frames lo .. hi inclusive)."""
    if world < 1 or not 0 <= rank < world or n_pairs < 0:
        raise ValueError("need world >= 1, 0 <= rank < world, n_pairs >= 0")
    q, r = divmod(n_pairs, world)
    lo = rank * q + min(rank, r)
    return lo, lo + q + (1 if rank < r else 0)


def shard_frames(n_frames: int, world: int, rank: int) -> Tuple[int, int]:
    """Frames [lo, hi) a rank needs for its share of the n_frames-1 consecutive pairs of a sequence
    (one overlapping frame between neighbouring ranks); (0, 0) if the rank has no pair."""
    lo, hi = shard_pairs(max(n_frames - 1, 0), world, rank)
    return (lo, hi + 1) if hi > lo else (0, 0)


STAT_KEYS = ("patches", "patch_iterations", "sampled_iterations", "resets", "start_oob", "kernel_launches")


def reduce_stats(stats: Dict[str, int], device_ms: float, dist=None, device=None) -> Dict[str, float]:
    """SUM of the run counters and MAX of the device time over all ranks (identity without `dist`)."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return {**{k: int(stats[k]) for k in STAT_KEYS}, "device_ms": float(device_ms)}
    import torch
    t = torch.tensor([float(stats[k]) for k in STAT_KEYS], dtype=torch.float64, device=device)
    m = torch.tensor([float(device_ms)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    dist.all_reduce(m, op=dist.ReduceOp.MAX)
    return {**{k: int(v) for k, v in zip(STAT_KEYS, t.tolist())}, "device_ms": float(m.item())}


def gather_results(local: List, dist=None) -> List:
    """Per-rank Python objects (e.g. per-pair EPE lists) gathered on every rank, in rank order."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return [local]
    out = [None] * dist.get_world_size()
    dist.all_gather_object(out, local)
    return out
