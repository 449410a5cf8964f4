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


@pytest.mark.parametrize("w,h,finest,coarsest", [(256, 112, 1, 3), (250, 107, 1, 3), (320, 200, 2, 3), (200, 100, 3, 3)])
def test_finest_scale_epilogue_equals_cv2_resize(oracle_port, w, h, finest, coarsest):
    """finest_scale > 0: the engine stops early, then flow * 2^f, a x2^f INTER_LINEAR upsample and the crop of the
    divisibility padding (src/main.cpp:192-198).  The level flow is parity-gated bit for bit against the oracle; the
    epilogue against OpenCV ITSELF (cv2.resize, optimised paths off), bit for bit — ragged sizes (crop offsets) and
    scales 2, 4, 8 included."""
    cv2 = pytest.importorskip("cv2")
    cv2.setUseOptimized(False)
    cv2.setNumThreads(1)
    a, b, _ = dis_b200.synth.make_pair(w, h, seed=44)
    e = dis_b200.Engine(0, finest_scale=finest, coarsest_scale=coarsest, max_iter=40)
    try:
        out = e.run_pair(a, b)
        lf = e.level_flow(0, 0)
    finally:
        e.close()
    _, d = oracle_port.run_pair_u8(a, b, finest=finest, coarsest=coarsest, iterations=40, dump=True)
    assert np.array_equal(bits(lf), bits(d["flow"][0]))
    sc = 1 << finest
    up = cv2.resize(lf * np.float32(sc), None, fx=sc, fy=sc, interpolation=cv2.INTER_LINEAR)    # main.cpp:193-195
    pw, ph = oracle_port.div_pad(w, coarsest), oracle_port.div_pad(h, coarsest)
    exp = up[ph // 2:ph // 2 + h, pw // 2:pw // 2 + w]                                           # main.cpp:198
    assert out.shape == (h, w, 2)
    assert np.array_equal(bits(out), bits(np.ascontiguousarray(exp)))


def test_colour_coding_matches_oracle(eng, oracle_port):
    """dis_color_code == draw_optical_flow (/root/reference/src/color_coding.cpp:83-117), byte for byte:
    automatic range (maxmotion <= 0), fixed range, invalid vectors, and the device-resident variant."""
    a, b, _ = dis_b200.synth.make_pair(256, 112, seed=9)
    flow = eng.run_pair(a, b)
    ref = oracle_port.color_code(flow)
    import oracle
    if oracle.load_refviz() is not None:     # the reference's own src/color_coding.cpp, compiled unmodified (oracle/_ref)
        rv = oracle.RefViz()
        assert np.array_equal(eng.color_code(flow), rv.color_code(flow))
        assert np.array_equal(eng.color_code(flow, 1.5), rv.color_code(flow, 1.5))
    assert np.array_equal(eng.color_code(flow), ref)
    assert np.array_equal(eng.color_code_pair(0, 112, 256), ref)
    assert np.array_equal(eng.color_code(flow, 1.5), oracle_port.color_code(flow, 1.5))
    rng = np.random.default_rng(3)
    f = (rng.normal(size=(97, 131, 2)) * 4).astype(np.float32)
    f[0, 0] = np.nan; f[5, 7] = (3e9, 1); f[9, 9] = 0; f[10, 10] = (0, -0.0); f[11, 11] = (-0.0, 0.0)
    assert np.array_equal(eng.color_code(f), oracle_port.color_code(f))
    z = np.zeros((8, 8, 2), np.float32)
    assert (eng.color_code(z) == 255).all()
