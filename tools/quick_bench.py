"""Development loop helper (GPU box): golden parity of two cases, then the bench workload's
per-level device times.  usage: python tools/quick_bench.py [reps] [n_frames]"""
import hashlib
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import dis_b200  # noqa: E402

z = np.load(os.path.join(ROOT, "tests", "golden", "golden.npz"))
meta = {m["name"]: m for m in json.loads(str(z["meta"]))}
for name in ("small_256x112", "sintel_1024x436"):
    m = meta[name]
    a, b, _ = dis_b200.synth.make_pair(m["w"], m["h"], seed=m["seed"])
    e = dis_b200.Engine(0)
    f = e.run_pair(a, b)
    e.close()
    ok = hashlib.sha256(f.tobytes()).hexdigest() == m["flow_sha"]
    print(f"parity {name}: {'bit-identical' if ok else 'MISMATCH'}", flush=True)

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 3
n_frames = int(sys.argv[2]) if len(sys.argv) > 2 else 50
frames, _ = dis_b200.synth.make_sequence(n_frames, 1024, 436, seed=100)
e = dis_b200.Engine(0)
first = np.arange(n_frames - 1, dtype=np.int32)
e.upload_frames(frames, first, first + 1)
e.run_resident(); e.synchronize()
e.timer_start()
for _ in range(reps):
    e.run_resident()
ms = e.timer_stop_ms() / reps
st = e.stats()
lv = [round(e.last_search_level_ms(l), 2) for l in range(e.num_levels())]
ph = [[round(e.last_search_phase_ms(l, p), 2) for l in range(e.num_levels())] for p in range(3)]
print(f"  converge {ph[0]} = {sum(ph[0]):.2f} | walk {ph[1]} = {sum(ph[1]):.2f} | stragglers {ph[2]} = {sum(ph[2]):.2f} ms")
print(f"{(n_frames - 1) / ms * 1e3:.1f} pairs/s  step {ms:.2f} ms  search levels fine->coarse {lv}  "
      f"samples/patch {st['sampled_iterations'] / st['patches']:.2f}")
e.close()
