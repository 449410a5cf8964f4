"""The host-buffer path behind bench.py's headline e2e number, at the sizes where it differs from the small-batch path:
the finest level runs in SLICES of pairs (>= 8 pairs; 3:1 for >= 32) whose flows go home on a copy stream while the
next slice is searched (csrc/dis_api.cu run_levels_group), and runs of more than 128 pairs are CHUNKED.  Every flow
must be bit-identical to the same pair run alone (dis_run_pair_u8), whatever the slicing.  Plus a stress test of the
walk <-> k_stragglers hand-off protocol: hundreds of back-to-back small runs, with and without CUDA_LAUNCH_BLOCKING."""
import os
import subprocess
import sys

import numpy as np
import pytest

import dis_b200

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def tune(name, value):
    assert dis_b200.load_library().dis_tune(name.encode(), None if value is None else str(value).encode()) == 0


@pytest.fixture(scope="module")
def seq():
    frames, _ = dis_b200.synth.make_sequence(41, 96, 64, seed=11)
    e = dis_b200.Engine(0)
    alone = np.stack([e.run_pair(frames[i], frames[i + 1]) for i in range(40)])
    e.close()
    return frames, alone


@pytest.mark.parametrize("n_pairs", [9, 40])
@pytest.mark.parametrize("slices", [None, "4,3,2,1", "1", "3", "7"])
def test_sliced_sequence_equals_single_pairs(seq, n_pairs, slices):
    frames, alone = seq
    tune("DIS_OUT_SLICES", slices)
    try:
        e = dis_b200.Engine(0)
        out = e.run_sequence(frames[:n_pairs + 1])
        again = e.run_sequence(frames[:n_pairs + 1])
        e.close()
    finally:
        tune("DIS_OUT_SLICES", None)
    assert np.array_equal(bits(out), bits(alone[:n_pairs]))
    assert np.array_equal(bits(out), bits(again))


def test_sliced_batch_equals_single_pairs(seq):
    frames, alone = seq
    e = dis_b200.Engine(0)
    out = e.run_batch(frames[:40], frames[1:41])
    e.close()
    assert np.array_equal(bits(out), bits(alone))


@pytest.mark.parametrize("dense_from", [1, 1000000])
def test_both_builds_of_k_stragglers_agree(seq, dense_from):
    """ps 8 has two builds of k_stragglers (two or three blocks per SM, csrc/dis_search.cu); launch_ps picks by the number
    of pairs in the launch (DIS_STRAG_DENSE).  Forced either way, a 40-pair run equals the pairs run alone."""
    frames, alone = seq
    tune("DIS_STRAG_DENSE", dense_from)
    try:
        e = dis_b200.Engine(0)
        first = np.arange(40, dtype=np.int32)
        e.upload_frames(frames, first, first + 1)
        e.run_resident()
        out = e.download_flows()
        one = e.run_pair(frames[3], frames[4])
        e.close()
    finally:
        tune("DIS_STRAG_DENSE", None)
    assert np.array_equal(bits(out), bits(alone))
    assert np.array_equal(bits(one), bits(alone[3]))


def test_chunked_sequence_of_130_frames():
    """> 128 pairs: dis_run_sequence_u8 works in chunks of 128 pairs; the last chunk here has one pair."""
    frames, _ = dis_b200.synth.make_sequence(131, 64, 48, seed=12)
    e = dis_b200.Engine(0)
    out = e.run_sequence(frames)
    st = e.stats()
    for i in (0, 63, 127, 128, 129):
        assert np.array_equal(bits(out[i]), bits(e.run_pair(frames[i], frames[i + 1]))), f"pair {i}"
    e.close()
    assert out.shape == (130, 48, 64, 2)
    assert st["patch_iterations"] > 0 and st["sampled_converge"] + st["sampled_walk"] <= st["sampled_iterations"]


STRESS = r"""
import sys, numpy as np
sys.path.insert(0, %r)
import dis_b200
sizes = [(64, 48), (96, 64), (128, 64), (160, 96), (200, 100), (256, 112)]
pairs = [dis_b200.synth.make_pair(w, h, seed=40 + i)[:2] for i, (w, h) in enumerate(sizes)]
e = dis_b200.Engine(0)
ref = [e.run_pair(a, b) for a, b in pairs]
bad = 0
for it in range(200):
    a, b = pairs[it %% len(pairs)]
    f = e.run_pair(a, b)          # raises on DIS_ERR_CUDA (the spin protocol's time-out flag)
    bad += not np.array_equal(f.view(np.uint32), ref[it %% len(pairs)].view(np.uint32))
e.close()
print("stress", "OK" if bad == 0 else "MISMATCH %%d" %% bad)
sys.exit(1 if bad else 0)
""" % ROOT


@pytest.mark.parametrize("blocking", ["0", "1"])
def test_stress_200_small_runs(blocking):
    env = dict(os.environ, CUDA_LAUNCH_BLOCKING=blocking)
    r = subprocess.run([sys.executable, "-c", STRESS], env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0 and "stress OK" in r.stdout, (r.stdout[-500:], r.stderr[-1500:])
