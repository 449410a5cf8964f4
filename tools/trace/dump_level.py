"""Writes the planes of each level of a synthetic pair (as the oracle builds them) for tools/trace/orbit_stats.c.
usage: python tools/trace/dump_level.py outprefix [seq|pair] [seed]"""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import dis_b200, oracle
pre = sys.argv[1]; kind = sys.argv[2] if len(sys.argv) > 2 else "seq"; seed = int(sys.argv[3]) if len(sys.argv) > 3 else 100
if kind == "seq":
    fr, _ = dis_b200.synth.make_sequence(3, 1024, 436, seed=seed); a, b = fr[0], fr[1]
else:
    a, b, _ = dis_b200.synth.make_pair(1024, 436, seed=seed)
o = oracle.Oracle("port")
fa, pw, ph = o.pad_convert(a, 3); fb, _, _ = o.pad_convert(b, 3)
H, W = fa.shape
pa = o.pyramid(fa, 4, 8); pb = o.pyramid(fb, 4, 8)
flow, d = o.optical_flow(pa, pb, 8, W, H, dump=True)
for s in range(4):
    with open(f"{pre}_L{s}.bin", "wb") as f:
        np.array([W, H, s, 8, 8, 2, 1000, 1, 1 if s < 3 else 0], np.int32).tofile(f)
        pa[1][s].tofile(f); pa[2][s].tofile(f); pb[0][s].tofile(f)
        if s < 3: d["flow"][s + 1].tofile(f)
print("ok")
