#!/usr/bin/env python
"""Experiment: do two engines on ONE GPU, each driven by its own host thread through the blocking
dis_run_sequence_u8, hide each other's host<->device copies?  (Whole steps alternate between the engines.)

  python tools/exp_two_engines.py [--steps 16] [--engines 2] [--slices 1]
Prints ms per step of: one engine (the bench's e2e loop), then E engines on E threads.
"""
import argparse
import os
import sys
import threading
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=16)
    ap.add_argument("--engines", type=int, default=2)
    ap.add_argument("--slices", default="")
    ap.add_argument("--stagger-ms", type=float, default=0.0)
    args = ap.parse_args()
    import torch
    import dis_b200
    import bench
    frames, _ = bench.bench_sequence(100)
    n, h, w = frames.shape
    engs, pins = [], []
    for _ in range(args.engines):
        e = dis_b200.Engine(0)
        pi = torch.empty((n, h, w), dtype=torch.uint8).pin_memory()
        po = torch.empty((n - 1, h, w, 2), dtype=torch.float32).pin_memory()
        pi.numpy()[:] = frames
        for _ in range(2):
            e.run_sequence_ptr(pi.data_ptr(), n, w, h, po.data_ptr())
        engs.append(e); pins.append((pi, po))
    ref = pins[0][1].numpy().copy()

    def loop(i, k, delay=0.0):
        if delay:
            time.sleep(delay)
        pi, po = pins[i]
        for _ in range(k):
            engs[i].run_sequence_ptr(pi.data_ptr(), n, w, h, po.data_ptr())

    t0 = time.perf_counter(); loop(0, args.steps); one = (time.perf_counter() - t0) / args.steps
    print(f"one engine, blocking calls: {1e3 * one:.2f} ms/step ({(n - 1) / one:.0f} pairs/s)", flush=True)
    for slices in [s for s in args.slices.split(";") if s] or [None]:
        if slices is not None:
            dis_b200.load_library().dis_tune(b"DIS_OUT_SLICES", slices.encode())
        for stagger in (0.0, args.stagger_ms * 1e-3) if args.stagger_ms else (0.0,):
            k = args.steps // args.engines
            th = [threading.Thread(target=loop, args=(i, k, stagger * i)) for i in range(args.engines)]
            t0 = time.perf_counter()
            for t in th:
                t.start()
            for t in th:
                t.join()
            dt = (time.perf_counter() - t0) / (k * args.engines)
            same = all(np.array_equal(p[1].numpy(), ref) for p in pins)
            print(f"{args.engines} engines on {args.engines} threads, slices={slices}, stagger {1e3 * stagger:.0f} ms: {1e3 * dt:.2f} ms/step "
                  f"({(n - 1) / dt:.0f} pairs/s), bit-identical outputs: {same}", flush=True)


if __name__ == "__main__":
    main()
