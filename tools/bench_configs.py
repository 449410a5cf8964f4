"""Throughput of the BASELINE.json configurations other than the headline one (informational; bench.py
measures configs[1]).  Device-resident inputs, CUDA-event timing on the engine's stream, exact mode.

  python tools/bench_configs.py                       # configs 3, 4, 5 on one GPU
  torchrun --nproc-per-node N tools/bench_configs.py  # adds config 5 row-tiled over N GPUs (strong scaling)
"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dis_b200  # noqa: E402

rank = int(os.environ.get("RANK", "0"))
local = int(os.environ.get("LOCAL_RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
out = {"n_gpus": world}


def resident(frames_a, frames_b, reps, **kw):
    n = len(frames_a)
    e = dis_b200.Engine(local, **kw)
    fr = np.concatenate([frames_a, frames_b])
    e.upload_frames(fr, np.arange(n, dtype=np.int32), np.arange(n, 2 * n, dtype=np.int32))
    e.run_resident(); e.synchronize()
    e.timer_start()
    for _ in range(reps):
        e.run_resident()
    ms = e.timer_stop_ms() / reps
    st = e.stats()
    e.close()
    return ms, st


if rank == 0:
    # config 3: a stream of independent 1024x436 pairs (here 64 resident pairs per pass)
    pairs = [dis_b200.synth.make_pair(1024, 436, seed=300 + i)[:2] for i in range(64)]
    ms, st = resident(np.stack([p[0] for p in pairs]), np.stack([p[1] for p in pairs]), 3)
    out["config3_batch_1024x436"] = {"pairs": 64, "ms_per_pass": ms, "pairs_per_s": 64 / ms * 1e3}
    # config 4: patch 12, overlap 0.75, scales 5 -> 0 at 1920x1080
    pairs = [dis_b200.synth.make_pair(1920, 1080, seed=400 + i, amp=5.6)[:2] for i in range(8)]
    ms, st = resident(np.stack([p[0] for p in pairs]), np.stack([p[1] for p in pairs]), 2,
                      patch_size=12, patch_overlap=0.75, coarsest_scale=5)
    out["config4_hq_1920x1080_patch12"] = {"pairs": 8, "ms_per_pass": ms, "pairs_per_s": 8 / ms * 1e3,
                                            "mpix_per_s": 8 * 1920 * 1080 / ms / 1e3, "patches": st["patches"]}

# config 5: one 3840x2160 pair, whole on one GPU and (world > 1) tiled by grid rows over all ranks
a, b, _ = dis_b200.synth.make_pair(3840, 2160, seed=500, amp=11.0)
if rank == 0:
    ms, st = resident(a[None], b[None], 2)
    out["config5_4k_one_gpu"] = {"ms_per_pair": ms, "pairs_per_s": 1e3 / ms, "mpix_per_s": 3840 * 2160 / ms / 1e3,
                                  "patches": st["patches"]}
if world > 1:
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    e = dis_b200.Engine(local)
    flow = dis_b200.parallel.run_pair_tiled(e, a, b, rank, world, dist)   # warm-up
    dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    reps = 3
    for _ in range(reps):
        flow = dis_b200.parallel.run_pair_tiled(e, a, b, rank, world, dist)
    torch.cuda.synchronize(); dist.barrier()
    dt = torch.tensor([(time.perf_counter() - t0) / reps], dtype=torch.float64, device="cuda")
    dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    ref = e.run_pair(a, b) if rank == 0 else None
    one = None
    if rank == 0:   # the same host-to-host call on one GPU, for a like-for-like comparison
        t1 = time.perf_counter()
        for _ in range(reps):
            e.run_pair(a, b)
        one = (time.perf_counter() - t1) / reps * 1e3
    e.close()
    if rank == 0:
        out["config5_4k_row_tiled"] = {"n_gpus": world, "ms_per_pair_host_to_host": float(dt.item()) * 1e3,
                                       "ms_per_pair_host_to_host_one_gpu": one,
                                       "note": "includes the H2D of the pair and the D2H of the 66 MB flow on every rank; wall clock, max over ranks",
                                       "bit_identical_to_one_gpu": bool(np.array_equal(flow.view(np.uint32), ref.view(np.uint32)))}
    dist.destroy_process_group()
if rank == 0:
    print(json.dumps(out))
