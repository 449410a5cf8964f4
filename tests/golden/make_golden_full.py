#!/usr/bin/env python
"""Generates tests/golden/golden_full.npz from the REFERENCE ITSELF (oracle/_ref) at the FULL sizes and
the FULL iteration budget (max_iter 1000) of every BASELINE.json config, plus the pairs bench.py times.

Run here (where /root/reference exists):  python tests/golden/make_golden_full.py [-j 8] [--only name,...]
One case per process (the reference is single-threaded; 3840x2160 takes ~15 min on one core).
Cases already present in golden_full.npz are kept unless named in --only.

Stored per case: sha256 of the input frames, of every level's dense flow, of every level's grid
coordinates and of the final flow; level geometry; a subsampled copy of the final flow; EPE vs the
synthetic ground truth.  tests/test_gpu_golden.py and bench.py compare the CUDA engine with these
WITHOUT running any oracle code.
"""
import argparse
import hashlib
import json
import multiprocessing as mp
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
OUT = os.path.join(HERE, "golden_full.npz")

# inputs: ("pair", w, h, seed, amp) -> synth.make_pair; ("seq", n, w, h, seed, i) -> pair i of synth.make_sequence;
#         ("lowtex", w, h, seed) -> synth.make_pair with the right half of both frames flattened (det == 0 patches)
CASES = [
    dict(name="sintel_seed1", src=("pair", 1024, 436, 1, None), params={}),
    dict(name="sintel_seed2", src=("pair", 1024, 436, 2, None), params={}),
    dict(name="sintel_amp14", src=("pair", 1024, 436, 3, 14.0), params={}),                 # large motion: many resets
    # half the frame flat: the compiled REFERENCE crashes on this input (a solve yields NaN and src/patch.cpp:222-234 then
    # indexes out of bounds), so this one golden comes from the C restatement, whose convention
    # (NaN => reset to the initial u) the engine shares; "oracle": "port" marks it
    dict(name="sintel_lowtex", src=("lowtex", 1024, 436, 4), params={}, oracle="port"),
    dict(name="bench_seq100_pair0", src=("seq", 50, 1024, 436, 100, 0), params={}),        # the timed workload of bench.py
    dict(name="bench_seq100_pair24", src=("seq", 50, 1024, 436, 100, 24), params={}),
    dict(name="bench_seq100_pair48", src=("seq", 50, 1024, 436, 100, 48), params={}),
    dict(name="hq_1920x1080_ps12", src=("pair", 1920, 1080, 0, 5.6),                        # BASELINE config 4
         params=dict(patch_size=12, overlap=0.75, coarsest=5)),
    dict(name="uhd_3840x2160", src=("pair", 3840, 2160, 1, 11.0), params={}),              # BASELINE config 5
]


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def make_input(src):
    import dis_b200
    kind = src[0]
    if kind == "pair":
        _, w, h, seed, amp = src
        return dis_b200.synth.make_pair(w, h, seed=seed, amp=amp)
    if kind == "seq":
        _, n, w, h, seed, i = src
        frames, gt = dis_b200.synth.make_sequence(n, w, h, seed=seed)
        return frames[i].copy(), frames[i + 1].copy(), gt
    if kind == "lowtex":
        _, w, h, seed = src
        return dis_b200.synth.make_lowtex_pair(w, h, seed=seed)
    raise ValueError(kind)


def run_case(c):
    import oracle
    t0 = time.time()
    ref = oracle.Oracle(c.get("oracle", "reference"))
    a, b, gt = make_input(tuple(c["src"]))
    flow, d = ref.run_pair_u8(a, b, dump=True, **c["params"])
    w = a.shape[1]
    sub = 16 if w >= 3000 else 8 if w >= 1024 else 4
    m = dict(name=c["name"], src=list(c["src"]), params=c["params"], oracle=c.get("oracle", "reference"), w=int(a.shape[1]), h=int(a.shape[0]),
             frames_sha=sha(np.stack([a, b])), flow_sha=sha(flow), level_flow_sha=[sha(f) for f in d["flow"]],
             level_ref_sha=[sha(r) for r in d["ref"]], level_info=d["info"], sub=sub,
             epe_vs_gt=float(np.linalg.norm(flow - gt, axis=-1).mean()), seconds=time.time() - t0)
    return m, flow[::sub, ::sub].copy()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("-j", type=int, default=min(8, os.cpu_count() or 1))
    ap.add_argument("--only", default="")
    args = ap.parse_args()
    meta, arrays = {}, {}
    if os.path.exists(OUT):
        z = np.load(OUT)
        for m in json.loads(str(z["meta"])):
            meta[m["name"]] = m
            arrays[m["name"] + "_sub"] = z[m["name"] + "_sub"]
    only = set(x for x in args.only.split(",") if x)
    todo = [c for c in CASES if (c["name"] in only if only else c["name"] not in meta)]
    todo.sort(key=lambda c: -c["src"][1] * c["src"][2] if c["src"][0] != "seq" else -c["src"][2] * c["src"][3])
    with mp.get_context("spawn").Pool(args.j) as pool:
        for m, sub in pool.imap_unordered(run_case, todo):
            meta[m["name"]] = m
            arrays[m["name"] + "_sub"] = sub
            print(m["name"], m["flow_sha"][:16], "EPE vs GT %.4f" % m["epe_vs_gt"], "%.0f s" % m["seconds"], flush=True)
            order = [c["name"] for c in CASES if c["name"] in meta]
            np.savez_compressed(OUT, meta=json.dumps([meta[n] for n in order]), **{n + "_sub": arrays[n + "_sub"] for n in order})


if __name__ == "__main__":
    main()
