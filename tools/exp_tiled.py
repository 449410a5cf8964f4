"""Development aid: where does the time of the row-banded 4K run go?  Times dis_pool_run_pair_tiled_u8 for several device
lists against one engine, and replays the same schedule through the engine-level calls to read each engine's per-level
search times.  usage: python tools/exp_tiled.py [max_iter]"""
import ctypes as C
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import dis_b200  # noqa: E402

mi = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
a, b, _ = dis_b200.synth.make_pair(3840, 2160, seed=500, amp=11.0)
pa, pb = torch.from_numpy(a).pin_memory(), torch.from_numpy(b).pin_memory()
out = torch.empty((2160, 3840, 2), dtype=torch.float32).pin_memory()
lib = dis_b200.load_library()

e = dis_b200.Engine(0, max_iter=mi)
e.upload_frames(np.stack([a, b]), np.array([0], np.int32), np.array([1], np.int32))
e.run_resident(); e.synchronize(); e.timer_start()
for _ in range(3):
    e.run_resident()
ms = e.timer_stop_ms() / 3
nl = e.num_levels()
print(f"one engine: {ms:.2f} ms device; search per level fine->coarse {[round(e.last_search_level_ms(l), 2) for l in range(nl)]}"
      f" phases {[[round(e.last_search_phase_ms(l, p), 2) for l in range(nl)] for p in range(3)]}", flush=True)
e.close()

ng = torch.cuda.device_count()
lists = [[0], [0, 0]] + ([[0, 1]] if ng >= 2 else []) + ([[0, 1, 2, 3]] if ng >= 4 else []) + ([list(range(8))] if ng >= 8 else [])
for devs in lists:
    p = dis_b200.Pool(devs, max_iter=mi)
    p.run_pair_tiled_ptr(pa.data_ptr(), pb.data_ptr(), 3840, 2160, out.data_ptr())
    t0 = time.perf_counter(); dev = []
    for _ in range(3):
        p.run_pair_tiled_ptr(pa.data_ptr(), pb.data_ptr(), 3840, 2160, out.data_ptr())
        dev.append(p.last_tiled_device_ms())
    print(f"pool {devs}: {1e3 * (time.perf_counter() - t0) / 3:.2f} ms host-to-host, {np.mean(dev):.2f} ms device (slowest GPU)", flush=True)
    p.close()

# the same schedule by hand on two engines, to read their per-level search times
for devs in ([0, 0], [0, 1] if ng >= 2 else None):
    if devs is None:
        continue
    E = [dis_b200.Engine(d, max_iter=mi) for d in devs]
    R = len(E)
    for rep in range(2):
        for x in E:
            x.tiled_begin(a, b)
        for level in range(nl - 1, -1, -1):
            li = E[0].level_info(level)
            bands = [dis_b200.parallel.shard_pairs(li["nph"], R, r) for r in range(R)]
            for r, x in enumerate(E):
                x.tiled_search(level, *bands[r])
            for r, x in enumerate(E):
                for q in range(R):
                    if q != r:   # whole neighbour band: more than the halo, same dependency structure
                        assert lib.dis_tiled_fetch_u_rows(x._h, level, E[q]._h, bands[q][0], bands[q][1]) == 0
                assert lib.dis_tiled_densify_rows(x._h, level, 0, li["height"]) == 0
        for x in E:
            assert lib.dis_tiled_end(x._h) == 0
    for r, x in enumerate(E):
        print(f"  by hand {devs} engine {r}: search per level fine->coarse {[round(x.last_search_level_ms(l), 2) for l in range(nl)]}"
              f" phases {[[round(x.last_search_phase_ms(l, p), 2) for l in range(nl)] for p in range(3)]}", flush=True)
    for x in E:
        x.close()
