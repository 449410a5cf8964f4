"""Development aid (GPU box): host-to-host throughput of dis_run_sequence_u8 on bench.py's workload
(pinned buffers, H2D + D2H inside the timed region), with a SHA-256 of the flows so that settings
can be compared for identity.  usage: [DIS_OUT_SLICES=...] python tools/e2e_bench.py [steps]"""
import hashlib
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import dis_b200  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 10
N, W, H = 50, 1024, 436
frames, _ = dis_b200.synth.make_sequence(N, W, H, seed=100)
eng = dis_b200.Engine(0)
pin_in = torch.empty((N, H, W), dtype=torch.uint8).pin_memory()
pin_out = torch.empty((N - 1, H, W, 2), dtype=torch.float32).pin_memory()
pin_in.numpy()[:] = frames
for _ in range(3):
    eng.run_sequence_ptr(pin_in.data_ptr(), N, W, H, pin_out.data_ptr())
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(steps):
    eng.run_sequence_ptr(pin_in.data_ptr(), N, W, H, pin_out.data_ptr())
dt = time.perf_counter() - t0
sha = hashlib.sha256(pin_out.numpy().tobytes()).hexdigest()[:16]
print(f"e2e {(N - 1) * steps / dt:.1f} pairs/s  {1e3 * dt / steps:.2f} ms/step  flows sha {sha}", flush=True)
