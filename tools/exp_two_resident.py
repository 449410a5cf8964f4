#!/usr/bin/env python
"""Experiment: E engines with the whole 49-pair bench sequence resident, stepping alternately (no copies).
usage: python tools/exp_two_resident.py [E ...]"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dis_b200  # noqa: E402

frames, _ = dis_b200.synth.make_sequence(50, 1024, 436, seed=100)
idx = np.arange(49, dtype=np.int32)
K = 12
for E in [int(x) for x in sys.argv[1:]] or [1, 2, 3]:
    engs = []
    for g in range(E):
        e = dis_b200.Engine(0)
        e.upload_frames(frames, idx, idx + 1)
        e.run_resident(); e.synchronize()
        engs.append(e)
    t0 = time.perf_counter()
    for i in range(K):
        engs[i % E].run_resident()
    for e in engs:
        e.synchronize()
    dt = (time.perf_counter() - t0) / K
    print(f"{E} engine(s), whole sequence each, alternating steps: {1e3 * dt:.2f} ms per step ({49 / dt:.0f} pairs/s)", flush=True)
    for e in engs:
        e.close()
