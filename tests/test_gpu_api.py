"""GPU tests of every C-ABI entry point of include/dis_b200.h against the oracle: the batch /
sequence / resident pipelines, the OpticalFlowClass contract on host pyramids (dis_run_pyramids),
construct_pyramide on the device, the PatchGrid-level API, the fast (conv_eps) extension and the
error behaviour."""
import numpy as np
import pytest

import dis_b200
from test_gpu_parity import bits, TOL_DENSE

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng():
    e = dis_b200.Engine(0)
    yield e
    e.close()


@pytest.fixture(scope="module")
def pairs():
    out = [dis_b200.synth.make_pair(200, 100, seed=20 + i) for i in range(5)]
    return np.stack([p[0] for p in out]), np.stack([p[1] for p in out])


@pytest.fixture(scope="module")
def pair_flows(oracle_port, pairs):
    a, b = pairs
    return np.stack([oracle_port.run_pair_u8(a[i], b[i]) for i in range(len(a))])


def test_batch_matches_single_pairs(eng, pairs, pair_flows):
    a, b = pairs
    out = eng.run_batch(a, b)
    assert np.array_equal(bits(out), bits(pair_flows))
    st = eng.stats()
    assert st["kernel_launches"] > 0 and st["patches"] > 0


def test_sequence_shares_frames(eng, oracle_port):
    frames, _ = dis_b200.synth.make_sequence(5, 200, 100, seed=30)
    out = eng.run_sequence(frames)
    assert out.shape == (4, 100, 200, 2)
    for i in range(4):
        ref = oracle_port.run_pair_u8(frames[i], frames[i + 1])
        assert np.array_equal(bits(out[i]), bits(ref)), f"pair {i}"


def test_resident_pipeline_and_timers(eng, pairs, pair_flows):
    a, b = pairs
    n = len(a)
    frames = np.concatenate([a, b])
    eng.upload_frames(frames, np.arange(n), np.arange(n) + n)
    eng.timer_start()
    eng.run_resident()
    ms = eng.timer_stop_ms()
    out = eng.download_flows()
    assert ms > 0
    assert np.array_equal(bits(out), bits(pair_flows))
    sms, nl = eng.last_search_ms()
    assert nl % eng.num_levels() == 0 and sms > 0
    # idempotent: a second run over the same resident frames gives the same bits
    eng.run_resident()
    assert np.array_equal(bits(eng.download_flows()), bits(pair_flows))


def test_construct_pyramid_matches_oracle(eng, oracle_port):
    a, _, _ = dis_b200.synth.make_pair(256, 112, seed=40)
    fa, _, _ = oracle_port.pad_convert(a, 3)
    P, Px, Py = eng.construct_pyramid(fa, 3, 8)
    Po, Pxo, Pyo = oracle_port.pyramid(fa, 4, 8)
    for s in range(4):
        assert P[s].shape == Po[s].shape
        assert np.array_equal(bits(P[s]), bits(Po[s])), f"image plane, scale {s}"
        assert np.array_equal(bits(Px[s]), bits(Pxo[s])), f"dx plane, scale {s}"
        assert np.array_equal(bits(Py[s]), bits(Pyo[s])), f"dy plane, scale {s}"


def test_planes_of_a_run_match_oracle(eng, oracle_port):
    a, b, _ = dis_b200.synth.make_pair(250, 107, seed=41)    # needs divisibility padding
    eng.run_pair(a, b)
    for frame, img in enumerate((a, b)):
        f, _, _ = oracle_port.pad_convert(img, 3)
        planes = oracle_port.pyramid(f, 4, 8)
        for s in range(4):
            for which in range(3):
                got = eng.plane(frame, s, which, planes[which][s].shape)
                assert np.array_equal(bits(got), bits(planes[which][s])), (frame, s, which)


@pytest.mark.parametrize("finest", [0, 1])
def test_run_pyramids_is_the_constructor_contract(oracle_port, finest):
    """dis_run_pyramids == OpticalFlowClass ctor (include/optical_flow.hpp:43-54) on host pyramids."""
    a, b, _ = dis_b200.synth.make_pair(256, 112, seed=42)
    fa, _, _ = oracle_port.pad_convert(a, 3)
    fb, _, _ = oracle_port.pad_convert(b, 3)
    H, W = fa.shape
    pa, pb = oracle_port.pyramid(fa, 4, 8), oracle_port.pyramid(fb, 4, 8)
    ref = oracle_port.optical_flow(pa, pb, 8, W, H, finest=finest)
    e = dis_b200.Engine(0, finest_scale=finest)
    try:
        out = e.run_pyramids(pa, pb, 8, W, H)
        assert out.shape == ref.shape == (H >> finest, W >> finest, 2)
        assert np.array_equal(bits(out), bits(ref))
        # the three pyramids the reference never reads may be omitted
        none = [None] * 4
        out2 = e.run_pyramids((none, pa[1], pa[2]), (pb[0], none, none), 8, W, H)
        assert np.array_equal(bits(out2), bits(ref))
    finally:
        e.close()


def test_img_padding_larger_than_patch_size(eng, oracle_port):
    a, b, _ = dis_b200.synth.make_pair(128, 64, seed=43)
    fa, _, _ = oracle_port.pad_convert(a, 3)
    fb, _, _ = oracle_port.pad_convert(b, 3)
    H, W = fa.shape
    pa, pb = oracle_port.pyramid(fa, 4, 11), oracle_port.pyramid(fb, 4, 11)
    ref = oracle_port.optical_flow(pa, pb, 11, W, H)
    assert np.array_equal(bits(eng.run_pyramids(pa, pb, 11, W, H)), bits(ref))


def test_finest_scale_epilogue(oracle_port):
    """finest_scale = 1: the engine stops one level early, then flow*2 and a x2 bilinear upsample
    (src/main.cpp:192-196).  The level flow is parity-gated bit for bit; the upsample against a
    float64 restatement of cv::resize INTER_LINEAR within 1e-5 px."""
    a, b, _ = dis_b200.synth.make_pair(256, 112, seed=44)
    e = dis_b200.Engine(0, finest_scale=1)
    try:
        out = e.run_pair(a, b)
        _, d = oracle_port.run_pair_u8(a, b, finest=1, dump=True)
        lf = e.level_flow(0, 0)
        assert np.array_equal(bits(lf), bits(d["flow"][0]))
        Hf, Wf = lf.shape[:2]
        ys = (np.arange(2 * Hf) + 0.5) / 2 - 0.5
        xs = (np.arange(2 * Wf) + 0.5) / 2 - 0.5
        y0 = np.clip(np.floor(ys).astype(int), 0, Hf - 1); y1 = np.clip(y0 + 1, 0, Hf - 1)
        x0 = np.clip(np.floor(xs).astype(int), 0, Wf - 1); x1 = np.clip(x0 + 1, 0, Wf - 1)
        ay = np.clip(ys - np.floor(ys), 0, 1) * (ys >= 0) * (ys <= Hf - 1)
        ax = np.clip(xs - np.floor(xs), 0, 1) * (xs >= 0) * (xs <= Wf - 1)
        f = lf.astype(np.float64) * 2
        top = f[y0][:, x0] * (1 - ax)[None, :, None] + f[y0][:, x1] * ax[None, :, None]
        bot = f[y1][:, x0] * (1 - ax)[None, :, None] + f[y1][:, x1] * ax[None, :, None]
        up = top * (1 - ay)[:, None, None] + bot * ay[:, None, None]
        exp = up                      # 256x112 needs no divisibility padding, so the crop is the identity
        assert out.shape == (112, 256, 2)
        assert np.abs(out - exp).max() < 1e-5
    finally:
        e.close()


def test_colour_coding_matches_oracle(eng, oracle_port):
    """dis_color_code == draw_optical_flow (/root/reference/src/color_coding.cpp:83-117), byte for byte:
    automatic range (maxmotion <= 0), fixed range, invalid vectors, and the device-resident variant."""
    a, b, _ = dis_b200.synth.make_pair(256, 112, seed=9)
    flow = eng.run_pair(a, b)
    ref = oracle_port.color_code(flow)
    assert np.array_equal(eng.color_code(flow), ref)
    assert np.array_equal(eng.color_code_pair(0, 112, 256), ref)
    assert np.array_equal(eng.color_code(flow, 1.5), oracle_port.color_code(flow, 1.5))
    rng = np.random.default_rng(3)
    f = (rng.normal(size=(97, 131, 2)) * 4).astype(np.float32)
    f[0, 0] = np.nan; f[5, 7] = (3e9, 1); f[9, 9] = 0; f[10, 10] = (0, -0.0); f[11, 11] = (-0.0, 0.0)
    assert np.array_equal(eng.color_code(f), oracle_port.color_code(f))
    z = np.zeros((8, 8, 2), np.float32)
    assert (eng.color_code(z) == 255).all()
