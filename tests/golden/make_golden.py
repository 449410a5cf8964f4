#!/usr/bin/env python
"""Generates tests/golden/golden.npz from the REFERENCE ITSELF (oracle/_ref = the reference's
unmodified src/patch.cpp, src/patch_grid.cpp, src/optical_flow.cpp compiled in place).

Run here (where /root/reference exists):  python tests/golden/make_golden.py
The reference has no golden vectors of its own; these are the pins this repo
creates: for each case the sha256 of every level's dense flow and of the final
flow as the reference computes them, a subsampled copy of the final flow, and the sha256 of the
input frames (so a drift of the synthetic generator is told apart from a parity failure).
The pre-processing feeding the reference engine is oracle/dis_oracle.c's restatement of the OpenCV
calls (checked against cv2 in tests/test_oracle.py).
"""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import dis_b200  # noqa: E402
import oracle  # noqa: E402

CASES = [
    dict(name="tiny_128x64", w=128, h=64, seed=0, params={}),
    dict(name="small_256x112", w=256, h=112, seed=1, params={}),
    dict(name="ragged_250x107", w=250, h=107, seed=3, params={}),
    dict(name="hq_480x272_ps12", w=480, h=272, seed=2, params=dict(patch_size=12, overlap=0.75, coarsest=4)),
    dict(name="maxiter16_256x112", w=256, h=112, seed=4, params=dict(iterations=16)),
    dict(name="nonorm_192x96", w=192, h=96, seed=5, params=dict(patnorm=False)),
    dict(name="sintel_1024x436", w=1024, h=436, seed=0, params={}),      # BASELINE config 1
]


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    ref = oracle.Oracle("reference")
    meta, arrays = [], {}
    for c in CASES:
        a, b, gt = dis_b200.synth.make_pair(c["w"], c["h"], seed=c["seed"])
        flow, d = ref.run_pair_u8(a, b, dump=True, **c["params"])
        sub = 8 if c["w"] >= 1024 else 4
        m = dict(c)
        m.update(frames_sha=sha(np.stack([a, b])), flow_sha=sha(flow), level_flow_sha=[sha(f) for f in d["flow"]],
                 level_ref_sha=[sha(r) for r in d["ref"]], level_info=d["info"], sub=sub,
                 epe_vs_gt=float(np.linalg.norm(flow - gt, axis=-1).mean()))
        arrays[c["name"] + "_sub"] = flow[::sub, ::sub].copy()
        meta.append(m)
        print(c["name"], m["flow_sha"][:16], "EPE vs GT %.4f" % m["epe_vs_gt"], flush=True)
    np.savez_compressed(os.path.join(HERE, "golden.npz"), meta=json.dumps(meta), **arrays)


if __name__ == "__main__":
    main()
