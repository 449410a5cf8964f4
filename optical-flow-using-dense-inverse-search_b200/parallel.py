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


STAT_KEYS = ("patches", "patch_iterations", "sampled_iterations", "resets", "start_oob", "kernel_launches",
             "sampled_converge", "sampled_walk")


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


# ---- one large frame tiled by grid rows over several GPUs (BASELINE.json config 5) --------
def band_rows(nph: int, world: int, rank: int) -> Tuple[int, int]:
    """Grid rows [row0, row1) of a level searched by `rank`: contiguous bands, heights differing by at most 1."""
    return shard_pairs(nph, world, rank)


def _merge_bands(parts, npw: int, nph: int, world: int):
    """parts[r]: the full-size u tensor [npw*nph, 2] of rank r (valid in its own band) -> merged tensor."""
    import torch
    out = torch.empty_like(parts[0])
    v = out.view(npw, nph, 2)
    for r in range(world):
        r0, r1 = band_rows(nph, world, r)
        v[:, r0:r1] = parts[r].view(npw, nph, 2)[:, r0:r1]
    return out


def run_pair_tiled(engine, a, b, rank: int, world: int, dist=None):
    """One frame pair split over `world` GPUs (one engine / process each): per level every rank searches the
    patches of its band of grid rows, the per-patch flows are all-gathered over NCCL (a few MB per level, no
    pixel data), and every rank densifies the whole level.  Returns the full flow on every rank, bit-identical
    to engine.run_pair(a, b).  `dist` = torch.distributed (initialised, NCCL) or None when world == 1."""
    import torch
    dev = torch.device("cuda", engine.device)
    engine.tiled_begin(a, b)
    for level in range(engine.num_levels() - 1, -1, -1):
        li = engine.level_info(level)
        r0, r1 = band_rows(li["nph"], world, rank)
        engine.tiled_search(level, r0, r1)
        if world > 1:
            mine = torch.empty((li["num_patches"], 2), dtype=torch.float32, device=dev)
            engine.tiled_get_u(level, mine.data_ptr())
            bh = -(-li["nph"] // world)                       # bands padded to one height for the all-gather
            send = torch.zeros((li["npw"], bh, 2), dtype=torch.float32, device=dev)
            send[:, : r1 - r0] = mine.view(li["npw"], li["nph"], 2)[:, r0:r1]
            recv = torch.empty((world, li["npw"], bh, 2), dtype=torch.float32, device=dev)
            dist.all_gather_into_tensor(recv, send)
            full = mine.view(li["npw"], li["nph"], 2)
            for r in range(world):
                q0, q1 = band_rows(li["nph"], world, r)
                full[:, q0:q1] = recv[r, :, : q1 - q0]
            torch.cuda.synchronize(dev)
            engine.tiled_set_u(level, mine.data_ptr())
        engine.tiled_densify(level)
    return engine.tiled_finish()


def run_pair_tiled_emulated(engines, a, b):
    """The same schedule with all `len(engines)` ranks living in ONE process (engines may share a GPU): the
    exchange is a device-side merge instead of an NCCL all-gather.  Used by the single-GPU parity test."""
    import torch
    world = len(engines)
    for e in engines:
        e.tiled_begin(a, b)
    for level in range(engines[0].num_levels() - 1, -1, -1):
        li = engines[0].level_info(level)
        parts = []
        for r, e in enumerate(engines):
            r0, r1 = band_rows(li["nph"], world, r)
            e.tiled_search(level, r0, r1)
            t = torch.empty((li["num_patches"], 2), dtype=torch.float32, device=torch.device("cuda", e.device))
            e.tiled_get_u(level, t.data_ptr())
            parts.append(t)
        for r, e in enumerate(engines):
            dev = torch.device("cuda", e.device)
            merged = _merge_bands([p.to(dev) for p in parts], li["npw"], li["nph"], world)
            torch.cuda.synchronize(dev)
            e.tiled_set_u(level, merged.data_ptr())
            e.tiled_densify(level)
    return [e.tiled_finish() for e in engines]
