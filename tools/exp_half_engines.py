#!/usr/bin/env python
"""Experiment: the 49 pairs of the bench step split over G engines (contiguous blocks), every engine stepping on its own
stream with no join between steps: do groups that drift out of phase beat one engine with all 49 pairs?
usage: python tools/exp_half_engines.py [G ...]"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dis_b200  # noqa: E402

frames, _ = dis_b200.synth.make_sequence(50, 1024, 436, seed=100)
K = 10
for G in [int(x) for x in sys.argv[1:]] or [1, 2, 3]:
    engs = []
    for g in range(G):
        lo, hi = dis_b200.parallel.shard_pairs(49, G, g)
        e = dis_b200.Engine(0)
        idx = np.arange(lo, hi, dtype=np.int32)
        e.upload_frames(frames, idx, idx + 1)
        e.run_resident(); e.synchronize()
        engs.append(e)
    t0 = time.perf_counter()
    for _ in range(K):
        for e in engs:
            e.run_resident()
    for e in engs:
        e.synchronize()
    dt = (time.perf_counter() - t0) / K
    print(f"{G} engine(s) sharing the 49 pairs: {1e3 * dt:.2f} ms per step ({49 / dt:.0f} pairs/s)", flush=True)
    for e in engs:
        e.close()
