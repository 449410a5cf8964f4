"""Workload for compute-sanitizer (memcheck / racecheck / synccheck / initcheck), small enough for the tools'
slow-down but through every kernel and the CONCURRENT walk <-> k_stragglers hand-off (default DIS_SIDE_BLOCKS):
  * a 3-frame 1024x436 sequence (2 pairs, pyramid reuse) — the bench workload's shape,
  * one 480x272 pair at patch 12 / overlap 0.75 / scales 4->0 (the non-tiled stragglers path),
  * one 128x64 pair (tiny levels: 1-block launches, y-offset-0 level),
  * a 9-pair 256x112 sequence through the host-buffer call (sliced finest level + copy stream),
  * colour coding and the finest_scale=1 epilogue.
Each result is compared with a second run (determinism); exit code 1 on any difference or error."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dis_b200

ok = True


def same(a, b, what):
    global ok
    eq = np.array_equal(np.ascontiguousarray(a).view(np.uint32), np.ascontiguousarray(b).view(np.uint32))
    print(f"{what}: {'deterministic' if eq else 'DIFFERS between two runs'}", flush=True)
    ok = ok and eq


fr, _ = dis_b200.synth.make_sequence(3, 1024, 436, seed=100)
e = dis_b200.Engine(0)
same(e.run_sequence(fr), e.run_sequence(fr), "sintel 3-frame sequence")
e.close()

a, b, _ = dis_b200.synth.make_pair(480, 272, seed=2)
e = dis_b200.Engine(0, patch_size=12, patch_overlap=0.75, coarsest_scale=4)
same(e.run_pair(a, b), e.run_pair(a, b), "480x272 patch 12")
e.close()

a, b, _ = dis_b200.synth.make_pair(128, 64, seed=0)
e = dis_b200.Engine(0)
f1 = e.run_pair(a, b)
same(f1, e.run_pair(a, b), "128x64")
bgr = e.color_code(f1)
e.close()

fr, _ = dis_b200.synth.make_sequence(10, 256, 112, seed=7)
e = dis_b200.Engine(0)
same(e.run_sequence(fr), e.run_sequence(fr), "256x112 10-frame sequence (sliced output)")
e.close()

a, b, _ = dis_b200.synth.make_pair(256, 112, seed=1)
e = dis_b200.Engine(0, finest_scale=1)
same(e.run_pair(a, b), e.run_pair(a, b), "finest_scale 1 epilogue")
e.close()
print("sanitize_run:", "OK" if ok else "FAILED", flush=True)
sys.exit(0 if ok else 1)
