"""Development aid (GPU box): one level in isolation — a sequence of small frames searched at scale 0 only, so that a
coarse level's launch size can be studied alone.  usage: python tools/level_probe.py W H [n_frames] [reps]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import dis_b200  # noqa: E402

W, H = int(sys.argv[1]), int(sys.argv[2])
n_frames = int(sys.argv[3]) if len(sys.argv) > 3 else 50
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 3
frames, _ = dis_b200.synth.make_sequence(n_frames, W, H, seed=100)
e = dis_b200.Engine(0, coarsest_scale=0, finest_scale=0)
first = np.arange(n_frames - 1, dtype=np.int32)
e.upload_frames(frames, first, first + 1)
e.run_resident(); e.synchronize()
e.timer_start()
for _ in range(reps):
    e.run_resident()
ms = e.timer_stop_ms() / reps
st = e.stats()
ph = [round(e.last_search_phase_ms(0, p), 3) for p in range(3)]
print(f"{W}x{H} x {n_frames - 1} pairs: {st['patches'] } patches, step {ms:.3f} ms | converge {ph[0]} walk {ph[1]} stragglers(exposed) {ph[2]} | "
      f"samples/patch {st['sampled_iterations'] / st['patches']:.2f} iterations/patch {st['patch_iterations'] / st['patches']:.1f}")
e.close()
