"""Is the tuning over-fitted to the bench sequence?  Times the same engine (exact mode, defaults, no knob changed) on other
content types at Sintel shape — device-resident (CUDA events) and host-to-host — and prints sampled passes per patch,
hand-offs and pairs/s for each next to the headline workload:
  seq100      the bench workload (49 consecutive pairs of one slowly moving texture)
  seeds64     64 independent pairs, 64 different textures (BASELINE config 3's content)
  amp14       32 pairs with 14 px motion amplitude (4.7x the default: many resets, long convergence)
  lowtex      32 pairs whose right 40 % is exactly flat and whose middle fades out (singular / ill-conditioned Hessians)
usage: python tools/bench_content.py [reps]   (one JSON line per content type)"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dis_b200  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
W, H = 1024, 436


def run(name, frames, first, second):
    import torch
    n = len(first)
    e = dis_b200.Engine(0)
    e.upload_frames(frames, np.asarray(first, np.int32), np.asarray(second, np.int32))
    e.run_resident(); e.synchronize()
    e.timer_start()
    for _ in range(reps):
        e.run_resident()
    ms = e.timer_stop_ms() / reps
    st = e.stats()
    nl = e.num_levels()
    ph = [[round(e.last_search_phase_ms(l, p), 2) for l in range(nl)] for p in range(3)]
    # host to host through the batch entry point (pinned buffers)
    a = torch.from_numpy(np.ascontiguousarray(frames[np.asarray(first)])).pin_memory()
    b = torch.from_numpy(np.ascontiguousarray(frames[np.asarray(second)])).pin_memory()
    out = torch.empty((n, H, W, 2), dtype=torch.float32).pin_memory()
    e.run_batch(a.numpy(), b.numpy(), out.numpy())
    t0 = time.perf_counter()
    for _ in range(reps):
        e.run_batch(a.numpy(), b.numpy(), out.numpy())
    e2e_ms = (time.perf_counter() - t0) / reps * 1e3
    e.close()
    smp = st["sampled_iterations"]
    print(json.dumps({"content": name, "pairs": n, "pairs_per_s_device": n / ms * 1e3, "pairs_per_s_host_to_host": n / e2e_ms * 1e3,
                      "ms_per_pass": ms, "sampled_passes_per_patch": smp / st["patches"],
                      "sampled_split": {"k_converge": st["sampled_converge"] / st["patches"], "walk": st["sampled_walk"] / st["patches"],
                                        "k_stragglers": (smp - st["sampled_converge"] - st["sampled_walk"]) / st["patches"]},
                      "resets_per_1000_patches": 1e3 * st["resets"] / st["patches"],
                      "phase_ms_fine_to_coarse": {"k_converge": ph[0], "walk": ph[1], "k_stragglers_exposed": ph[2]}}), flush=True)


frames, _ = dis_b200.synth.make_sequence(50, W, H, seed=100)
run("seq100 (bench workload)", frames, np.arange(49), np.arange(49) + 1)
prs = [dis_b200.synth.make_pair(W, H, seed=300 + i)[:2] for i in range(64)]
fr = np.concatenate([np.stack([p[0] for p in prs]), np.stack([p[1] for p in prs])])
run("seeds64 (config 3 content)", fr, np.arange(64), np.arange(64) + 64)
prs = [dis_b200.synth.make_pair(W, H, seed=700 + i, amp=14.0)[:2] for i in range(32)]
fr = np.concatenate([np.stack([p[0] for p in prs]), np.stack([p[1] for p in prs])])
run("amp14 (large motion)", fr, np.arange(32), np.arange(32) + 32)
prs = [dis_b200.synth.make_lowtex_pair(W, H, seed=800 + i)[:2] for i in range(32)]
fr = np.concatenate([np.stack([p[0] for p in prs]), np.stack([p[1] for p in prs])])
run("lowtex (40 % flat)", fr, np.arange(32), np.arange(32) + 32)
