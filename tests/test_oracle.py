"""CPU tests that PIN the oracle (no GPU needed).

 * oracle/dis_oracle.c (the C restatement) must reproduce, bit for bit, the reference's own engine
   sources compiled in place (oracle/_ref) — where /root/reference exists — and the committed golden
   vectors in tests/golden/golden.npz, which were generated from that compiled reference.
 * its pre-processing half (OpenCV calls of src/main.cpp:12-50,139-160) is checked against cv2.
 * level geometry and the Eigen reduction tree have known-answer tests.
"""
import ctypes as C
import hashlib
import json
import os

import numpy as np
import pytest

import dis_b200

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden.npz")


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def golden_cases():
    z = np.load(GOLDEN)
    return json.loads(str(z["meta"])), z


META, ARR = golden_cases()
FAST_CASES = [m for m in META if m["w"] * m["h"] <= 256 * 112]


@pytest.mark.parametrize("m", FAST_CASES, ids=[m["name"] for m in FAST_CASES])
def test_port_matches_golden(oracle_port, m):
    a, b, _ = dis_b200.synth.make_pair(m["w"], m["h"], seed=m["seed"])
    if sha(np.stack([a, b])) != m["frames_sha"]:
        pytest.skip("synthetic generator drifted (numpy/scipy version): golden inputs not reproducible here")
    flow, d = oracle_port.run_pair_u8(a, b, dump=True, **m["params"])
    assert [sha(f) for f in d["flow"]] == m["level_flow_sha"]
    assert [sha(r) for r in d["ref"]] == m["level_ref_sha"]
    assert d["info"] == m["level_info"]
    assert sha(flow) == m["flow_sha"]
    assert np.array_equal(bits(flow[::m["sub"], ::m["sub"]]), bits(ARR[m["name"] + "_sub"]))


@pytest.mark.parametrize("w,h,seed,kw", [
    (128, 64, 10, {}),
    (200, 100, 11, {}),
    (160, 96, 12, dict(patch_size=12, overlap=0.75, coarsest=2)),
    (96, 64, 13, dict(patch_size=4, overlap=0.5, coarsest=2)),
    (96, 64, 14, dict(patch_size=6, overlap=0.7, coarsest=1)),
    (128, 64, 15, dict(patnorm=False)),
    (128, 64, 16, dict(iterations=3)),
    (128, 64, 17, dict(finest=1)),
])
def test_port_matches_compiled_reference(oracle_port, oracle_ref, w, h, seed, kw):
    """Bit-exact agreement of the restatement with the reference's own compiled sources, level by level."""
    a, b, _ = dis_b200.synth.make_pair(w, h, seed=seed, amp=4.0)
    fp, dp = oracle_port.run_pair_u8(a, b, dump=True, **kw)
    fr, dr = oracle_ref.run_pair_u8(a, b, dump=True, **kw)
    assert dp["info"] == dr["info"]
    for s in range(len(dp["flow"])):
        assert np.array_equal(dp["ref"][s], dr["ref"][s])
        assert np.array_equal(bits(dp["flow"][s]), bits(dr["flow"][s])), f"level {s}"
        # Patch::get_patch_pos() = ref + u (include/patch.hpp:41): the reference exposes u only this way
        q = dp["ref"][s] + dp["u"][s]
        assert np.array_equal(bits(q), bits(dr["query"][s])), f"level {s} query positions"
    assert np.array_equal(bits(fp), bits(fr))


def test_constructor_and_patchgrid_drivers_of_the_reference_agree(oracle_ref):
    """oracle/_ref through the OpticalFlowClass constructor == through PatchGrid's public API."""
    a, b, _ = dis_b200.synth.make_pair(128, 64, seed=18)
    f1 = oracle_ref.run_pair_u8(a, b)
    f2, _ = oracle_ref.run_pair_u8(a, b, dump=True)
    assert np.array_equal(bits(f1), bits(f2))


def test_preprocessing_matches_opencv(oracle_port):
    cv2 = pytest.importorskip("cv2")
    cv2.setUseOptimized(False)     # the AVX resize path differs by 1 ulp in 0.3 % of pixels
    a, _, _ = dis_b200.synth.make_pair(250, 107, seed=19)
    f, pw, ph = oracle_port.pad_convert(a, 3)
    ref = cv2.copyMakeBorder(a, ph // 2, ph - ph // 2, pw // 2, pw - pw // 2, cv2.BORDER_REPLICATE).astype(np.float32)
    assert np.array_equal(f, ref)                                             # src/main.cpp:139-160
    gx, gy = oracle_port.sobel(f)
    assert np.array_equal(bits(gx), bits(cv2.Sobel(f, cv2.CV_32F, 1, 0, ksize=3, scale=1 / 8.0)))   # :19,34
    assert np.array_equal(bits(gy), bits(cv2.Sobel(f, cv2.CV_32F, 0, 1, ksize=3, scale=1 / 8.0)))   # :20,35
    P, Px, Py = oracle_port.pyramid(f, 4, 8)
    lvl = cv2.sqrt(gx * gx + gy * gy)                                         # :21-26
    for i in range(4):
        if i:
            lvl = cv2.resize(lvl, None, fx=0.5, fy=0.5, interpolation=cv2.INTER_LINEAR)   # :29
        dx = cv2.Sobel(lvl, cv2.CV_32F, 1, 0, ksize=3, scale=1 / 8.0)
        dy = cv2.Sobel(lvl, cv2.CV_32F, 0, 1, ksize=3, scale=1 / 8.0)
        assert np.array_equal(bits(P[i]), bits(cv2.copyMakeBorder(lvl, 8, 8, 8, 8, cv2.BORDER_REPLICATE))), i   # :43
        assert np.array_equal(bits(Px[i]), bits(cv2.copyMakeBorder(dx, 8, 8, 8, 8, cv2.BORDER_CONSTANT, value=0))), i
        assert np.array_equal(bits(Py[i]), bits(cv2.copyMakeBorder(dy, 8, 8, 8, 8, cv2.BORDER_CONSTANT, value=0))), i


def test_level_geometry_table(oracle_port):
    """SURVEY.md §8 size table (src/optical_flow.cpp:38,51-59; src/patch_grid.cpp:20-25)."""
    assert oracle_port.port.oracle_steps(8, C.c_float(0.7)) == 2
    assert oracle_port.port.oracle_steps(12, C.c_float(0.75)) == 3
    assert oracle_port.port.oracle_steps(8, C.c_float(0.95)) == 1
    assert oracle_port.div_pad(436, 3) == 4 and oracle_port.div_pad(1024, 3) == 0 and oracle_port.div_pad(1080, 5) == 8
    exp = {3: (128, 55, 64, 28, 1, 0), 2: (256, 110, 128, 55, 1, 1), 1: (512, 220, 256, 110, 1, 1),
           0: (1024, 440, 512, 220, 1, 1)}
    tot = 0
    for s, (w, h, npw, nph, ow, oh) in exp.items():
        li = oracle_port.level_info(1024, 440, s, 8, 8, 0.7)
        assert (li["w"], li["h"], li["npw"], li["nph"], li["offw"], li["offh"]) == (w, h, npw, nph, ow, oh)
        assert li["tmp_w"] == w + 16 and li["steps"] == 2
        tot += npw * nph
    assert tot == 149632
    hq = [oracle_port.level_info(1920, 1088, s, 12, 12, 0.75) for s in range(6)]
    assert sum(i["npw"] * i["nph"] for i in hq) == 309960
    assert [i["offh"] for i in hq] == [1, 0, 1, 0, 1, 0]


def test_flat_patch_keeps_init(oracle_port):
    """det == 0 branch (src/patch.cpp:86-90): constant frames => zero flow, every patch runs max_iter+1 passes."""
    a = np.full((64, 96), 9, np.uint8)
    flow, d = oracle_port.run_pair_u8(a, a, dump=True, iterations=5)
    assert not flow.any()
    assert all((it == 6).all() for it in d["iters"])


def test_colour_coding_known_answers(oracle_port):
    """compute_color (/root/reference/src/color_coding.cpp:13-81): zero flow is white, unit flow along +x
    is wheel entry 0 (pure red, stored B,G,R), invalid vectors stay black, range > 1 is dimmed by 0.75."""
    f = np.zeros((2, 4, 2), np.float32)
    f[0, 1] = (1, 0); f[0, 2] = (np.nan, 0); f[0, 3] = (2e9, 0); f[1, 0] = (-1, 0); f[1, 1] = (0, 1)
    c = oracle_port.color_code(f)
    assert c[0, 0].tolist() == [255, 255, 255]
    assert c[0, 1].tolist() == [0, 0, 255]
    assert c[0, 2].tolist() == [0, 0, 0] and c[0, 3].tolist() == [0, 0, 0]
    assert c[1, 0].tolist() == [255, 255, 0] or c[1, 0][2] == 0          # opposite direction: cyan side of the wheel
    d = oracle_port.color_code(f, maxmotion=0.5)                          # radius 2 > 1: colour * 0.75
    assert d[0, 1].tolist() == [0, 0, 191]


# ---- N1 / N2 pinned against the reference ITSELF: src/color_coding.cpp and src/IO_flow.cpp compiled unmodified where
#      they lie against a functional miniature of the cv types they use (oracle/stubs_cv, oracle/_ref/libdis_refviz.so) ----
@pytest.fixture(scope="module")
def refviz():
    import oracle
    if oracle.load_refviz() is None:
        pytest.skip("oracle/_ref/libdis_refviz.so not built (no /root/reference here)")
    return oracle.RefViz()


def viz_fields(oracle_port):
    rng = np.random.default_rng(0)
    out = [(rng.standard_normal((h, w, 2)) * s).astype(np.float32) for w, h, s in ((64, 48, 1.0), (200, 100, 3.0), (333, 77, 0.2), (128, 64, 30.0))]
    a, b, gt = dis_b200.synth.make_pair(256, 112, seed=1)
    out += [oracle_port.run_pair_u8(a, b, iterations=30), gt.copy()]
    f = out[0].copy(); f[3, 4, 0] = np.nan; f[5, 6, 1] = 2e9; f[0, 0] = 0; f[7, 7] = (np.inf, 1)
    out.append(f)
    out.append(np.zeros((5, 9, 2), np.float32))
    return out


def test_colour_coding_port_equals_the_reference(oracle_port, refviz):
    """oracle_color_code (the restatement the GPU kernel is tested against) == draw_optical_flow of the reference's own
    src/color_coding.cpp:83-117, byte for byte, for automatic and fixed ranges, invalid vectors included."""
    for f in viz_fields(oracle_port):
        for mm in (-1.0, 0.0, 2.5):
            assert np.array_equal(oracle_port.color_code(f, mm), refviz.color_code(f, mm)), (f.shape, mm)


def test_flo_files_equal_the_reference(oracle_port, refviz, tmp_path):
    """dis_write_flo writes the bytes SaveFlowFile writes (src/IO_flow.cpp:58-98); each side reads the other's file."""
    for i, f in enumerate(viz_fields(oracle_port)[:6]):
        h, w, _ = f.shape
        mine, theirs = tmp_path / f"m{i}.flo", tmp_path / f"r{i}.flo"
        dis_b200.write_flo(str(mine), f)
        refviz.save_flo(theirs, f)
        assert open(mine, "rb").read() == open(theirs, "rb").read()
        assert np.array_equal(bits(dis_b200.read_flo(str(theirs))), bits(f))
        assert np.array_equal(bits(refviz.read_flo(mine, w, h)), bits(f))
