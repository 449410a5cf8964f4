"""Short profiling workload: the bench's 49-pair sequence, device-resident, `reps` passes."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dis_b200

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 2
n_frames = int(sys.argv[2]) if len(sys.argv) > 2 else 50
frames, _ = dis_b200.synth.make_sequence(n_frames, 1024, 436, seed=100)
e = dis_b200.Engine(0)
first = np.arange(n_frames - 1, dtype=np.int32)
e.upload_frames(frames, first, first + 1)
for _ in range(reps):
    e.run_resident()
e.synchronize()
print("stats", e.stats(), "search ms", e.last_search_ms())
e.close()
