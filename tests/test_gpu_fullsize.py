"""BASELINE.json configs 3-5 at their FULL sizes: bit parity against the oracle with an iteration budget
the CPU finishes in seconds, plus size-independent properties at the full 1000-iteration budget
(batch == individual pairs, run-to-run determinism, accuracy against the known synthetic flow)."""
import numpy as np
import pytest

import dis_b200

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def test_hq_1920x1080_patch12_matches_oracle(oracle_port):
    """config 4: patch 12, overlap 0.75, scales 5 -> 0 at 1920x1080 (padded to 1920x1088, 309 960 patches)."""
    a, b, gt = dis_b200.synth.make_pair(1920, 1080, seed=0, amp=5.6)
    kw = dict(patch_size=12, patch_overlap=0.75, coarsest_scale=5)
    e = dis_b200.Engine(0, max_iter=60, **kw)
    try:
        flow = e.run_pair(a, b)
        assert [e.level_info(s)["num_patches"] for s in range(6)] == [232320, 58240, 14560, 3680, 920, 240]
    finally:
        e.close()
    ref = oracle_port.run_pair_u8(a, b, iterations=60, patch_size=12, coarsest=5, overlap=0.75)
    assert np.array_equal(bits(flow), bits(ref))
    e = dis_b200.Engine(0, **kw)                      # the full budget: properties only
    try:
        f1 = e.run_pair(a, b)
        f2 = e.run_pair(a, b)
        st = e.stats()
    finally:
        e.close()
    assert np.array_equal(bits(f1), bits(f2))
    assert st["patches"] == 309960 and st["patch_iterations"] > 0.95 * 1001 * 309960
    assert np.linalg.norm(f1 - gt, axis=-1).mean() < 0.35


def test_4k_3840x2160_matches_oracle(oracle_port):
    """config 5 geometry on one GPU: 2 754 000 patches (the row-tiled multi-GPU variants — NCCL all-gather between processes,
    halo exchange over peer memory inside one process — are covered by tests/test_gpu_tiled.py)."""
    a, b, _ = dis_b200.synth.make_pair(3840, 2160, seed=1, amp=11.0)
    e = dis_b200.Engine(0, max_iter=26)
    try:
        flow = e.run_pair(a, b)
        assert sum(e.level_info(s)["num_patches"] for s in range(4)) == 2754000
        assert e.level_info(0)["npw"] == 1920 and e.level_info(0)["nph"] == 1080
    finally:
        e.close()
    ref = oracle_port.run_pair_u8(a, b, iterations=26)
    assert np.array_equal(bits(flow), bits(ref))


def test_batch_of_sintel_pairs_equals_single_pairs():
    """config 3 (a stream of independent 1024x436 pairs): a batch is bit-identical to its pairs run alone,
    whatever the batch composition — every patch's arithmetic is independent of its neighbours in the launch."""
    pairs = [dis_b200.synth.make_pair(1024, 436, seed=200 + i)[:2] for i in range(6)]
    a = np.stack([p[0] for p in pairs]); b = np.stack([p[1] for p in pairs])
    e = dis_b200.Engine(0)
    try:
        batch = e.run_batch(a, b)
        for i in (0, 3, 5):
            assert np.array_equal(bits(e.run_pair(a[i], b[i])), bits(batch[i]))
        rev = e.run_batch(a[::-1].copy(), b[::-1].copy())
        assert np.array_equal(bits(rev[::-1]), bits(batch))
    finally:
        e.close()
