"""GPU vs the committed golden vectors (tests/golden/golden.npz, golden_full.npz), which were generated from the
reference's own compiled sources by tests/golden/make_golden.py / make_golden_full.py.  No oracle code runs
here: the CUDA engine's output is hashed and compared with what the reference produced — at the full sizes and
the full 1000-iteration budget of every BASELINE.json config (1024x436 defaults incl. large motion, 1920x1080
patch 12 / overlap 0.75 / scales 5->0, 3840x2160) and for pairs 0, 24, 48 of the sequence bench.py times."""
import functools
import hashlib
import json
import os

import numpy as np
import pytest

import dis_b200

pytestmark = pytest.mark.gpu
GDIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
Z = np.load(os.path.join(GDIR, "golden.npz"))
META = json.loads(str(Z["meta"]))
# full sizes at the full iteration budget: every BASELINE.json config + the pairs bench.py times (make_golden_full.py)
ZF = np.load(os.path.join(GDIR, "golden_full.npz"))
META_FULL = json.loads(str(ZF["meta"]))
KEYS = {"iterations": "max_iter", "patch_size": "patch_size", "coarsest": "coarsest_scale", "finest": "finest_scale",
        "overlap": "patch_overlap", "patnorm": "patch_norm"}


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def golden_input(m):
    """The frames a golden case was generated from (tests/golden/make_golden*.py)."""
    src = m.get("src")
    if src is None:
        return dis_b200.synth.make_pair(m["w"], m["h"], seed=m["seed"])
    if src[0] == "pair":
        return dis_b200.synth.make_pair(src[1], src[2], seed=src[3], amp=src[4])
    if src[0] == "lowtex":
        return dis_b200.synth.make_lowtex_pair(src[1], src[2], seed=src[3])
    frames, gt = _sequence(src[1], src[2], src[3], src[4])
    return frames[src[5]].copy(), frames[src[5] + 1].copy(), gt


@functools.lru_cache(maxsize=2)
def _sequence(n, w, h, seed):
    return dis_b200.synth.make_sequence(n, w, h, seed=seed)


@pytest.mark.parametrize("m", META + META_FULL, ids=[m["name"] for m in META + META_FULL])
def test_engine_reproduces_reference_golden(m):
    a, b, gt = golden_input(m)
    Z = globals()["ZF"] if "src" in m else globals()["Z"]
    if sha(np.stack([a, b])) != m["frames_sha"]:
        pytest.skip("synthetic generator drifted (numpy/scipy version): golden inputs not reproducible here")
    e = dis_b200.Engine(0, **{KEYS[k]: (int(v) if isinstance(v, bool) else v) for k, v in m["params"].items()})
    try:
        flow = e.run_pair(a, b)
        nl = e.num_levels()
        assert nl == len(m["level_info"])
        for s in range(nl):
            li = e.level_info(s)
            g = m["level_info"][s]
            assert (li["width"], li["height"], li["npw"], li["nph"], li["offw"], li["offh"], li["steps"], li["tmp_w"]) == \
                   (g["w"], g["h"], g["npw"], g["nph"], g["offw"], g["offh"], g["steps"], g["tmp_w"])
            assert sha(e.patches(0, s)[0]) == m["level_ref_sha"][s], f"grid coordinates, level {s}"
            assert sha(e.level_flow(0, s)) == m["level_flow_sha"][s], f"dense flow, level {s}"
        sub = flow[::m["sub"], ::m["sub"]]
        d = np.linalg.norm(sub - Z[m["name"] + "_sub"], axis=-1)
        assert d.mean() <= 0.01, "north-star dense tolerance"      # 0.01 px (BASELINE.json)
        assert sha(flow) == m["flow_sha"], f"not bit-identical to the reference (subsampled max diff {d.max()})"
        assert abs(float(np.linalg.norm(flow - gt, axis=-1).mean()) - m["epe_vs_gt"]) < 1e-6    # EPE delta vs ref
    finally:
        e.close()
