"""GPU tests of the PatchGrid-level entry points (include/patch_grid.hpp:16-30 mirrored by
dis_grid_*), known-answer patches, the conv_eps extension and error behaviour."""
import ctypes as C

import numpy as np
import pytest

import dis_b200
from dis_b200.engine import _chk, _fp
from test_gpu_parity import bits, TOL_DENSE

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng():
    e = dis_b200.Engine(0)
    yield e
    e.close()


class Grid:
    """Thin test-side wrapper of dis_grid_* (one level of one pair, host pointers)."""

    def __init__(self, eng, w, h, pad):
        self.lib, self.w, self.h = eng.lib, w, h
        self.g = C.c_void_p()
        _chk(self.lib.dis_grid_create(eng._h, w, h, pad, C.byref(self.g)))
        self.n = self.lib.dis_grid_num_patches(self.g)

    def close(self):
        self.lib.dis_grid_destroy(self.g)

    def init(self, img, dx, dy):
        _chk(self.lib.dis_grid_init(self.g, img.ctypes.data_as(_fp), dx.ctypes.data_as(_fp), dy.ctypes.data_as(_fp)))

    def target(self, img, dx, dy):
        _chk(self.lib.dis_grid_set_target(self.g, img.ctypes.data_as(_fp), dx.ctypes.data_as(_fp), dy.ctypes.data_as(_fp)))

    def from_prev(self, flow):
        flow = np.ascontiguousarray(flow, np.float32)
        _chk(self.lib.dis_grid_init_from_prev_flow(self.g, flow.ctypes.data_as(_fp)))

    def set_init(self, i, ux, uy):
        _chk(self.lib.dis_grid_set_patch_init(self.g, i, C.c_float(ux), C.c_float(uy)))

    def search(self):
        _chk(self.lib.dis_grid_inverse_search(self.g))

    def densify(self):
        out = np.empty((self.h, self.w, 2), np.float32)
        _chk(self.lib.dis_grid_densify(self.g, out.ctypes.data_as(_fp)))
        return out

    def patch(self, i):
        r, u = (C.c_float * 2)(), (C.c_float * 2)()
        _chk(self.lib.dis_grid_get_patch(self.g, i, r, u))
        return (r[0], r[1]), (u[0], u[1])


def test_grid_api_reproduces_the_level_loop(eng, oracle_port):
    """Drive dis_grid_* exactly as src/optical_flow.cpp:67-90 drives PatchGrid, coarse to fine."""
    a, b, _ = dis_b200.synth.make_pair(256, 112, seed=50)
    fa, _, _ = oracle_port.pad_convert(a, 3)
    fb, _, _ = oracle_port.pad_convert(b, 3)
    H, W = fa.shape
    pa, pb = oracle_port.pyramid(fa, 4, 8), oracle_port.pyramid(fb, 4, 8)
    ref, d = oracle_port.optical_flow(pa, pb, 8, W, H, dump=True)
    prev = None
    for scale in (3, 2, 1, 0):
        g = Grid(eng, W >> scale, H >> scale, 8)
        try:
            assert g.n == d["info"][scale]["npw"] * d["info"][scale]["nph"]
            g.init(pa[0][scale], pa[1][scale], pa[2][scale])
            g.target(pb[0][scale], pb[1][scale], pb[2][scale])
            if prev is not None:
                g.from_prev(prev)
            g.search()
            prev = g.densify()
            assert np.array_equal(bits(prev), bits(d["flow"][scale])), f"scale {scale}"
            for i in (0, g.n // 2, g.n - 1):
                r, u = g.patch(i)
                assert r == tuple(d["ref"][scale][i]) and np.array_equal(bits(np.array(u)), bits(d["u"][scale][i]))
        finally:
            g.close()
    assert np.array_equal(bits(prev), bits(ref))


def _one_level(oracle_port, seed=51, w=128, h=64):
    a, b, _ = dis_b200.synth.make_pair(w, h, seed=seed)
    fa = a.astype(np.float32); fb = b.astype(np.float32)
    pa, pb = oracle_port.pyramid(fa, 1, 8), oracle_port.pyramid(fb, 1, 8)
    return pa, pb


def test_known_answer_patches(eng, oracle_port):
    """Explicit per-patch inits (Patch::inverse_search(init), src/patch.cpp:156-162):
    a start outside the bounds keeps u = init with 0 passes (:131-138); a start near the bound that
    walks out is reset to init (:185-194); integer and just-below-integer positions exercise the
    ceil(p + 1e-5f) tap base (:222-234)."""
    w, h = 320, 64      # refs up to x = 319: p + 1e-5f == p in fp32 for p >= 256
    pa, pb = _one_level(oracle_port, 52, w, h)
    g = Grid(eng, w, h, 8)
    try:
        g.init(pa[0][0], pa[1][0], pa[2][0]); g.target(pb[0][0], pb[1][0], pb[2][0])
        info = oracle_port.level_info(w, h, 0, 8, 8, 0.7)
        nph = info["nph"]
        cases = {0: (-30.0, 0.0),                       # start out of bounds (x < lb)
                 1: (0.0, -9.5),                        # start out of bounds (y < lb)
                 nph * 140 + 10: (0.0, 0.0),            # ref.x = 281 >= 256: integer position quirk
                 nph * 140 + 11: (-0.000005, 0.0),      # just below an integer
                 nph * 20 + 5: (2.999995, -1.999995),
                 nph * 159 + 3: (3.5, 0.0),             # start in bounds near the right edge
                 g.n - 1: (400.0, 400.0)}
        for i, (ux, uy) in cases.items():
            g.set_init(i, ux, uy)
        g.search()
        # patches are independent within a level: untouched patches must equal the oracle's zero-init
        # run bit for bit; the special ones are checked against their defining properties
        _, d = oracle_port.optical_flow(pa, pb, 8, w, h, coarsest=0, dump=True)
        for i in range(0, g.n, 97):
            if i in cases:
                continue
            _, u = g.patch(i)
            assert np.array_equal(bits(np.array(u)), bits(d["u"][0][i])), i
        for i in (0, 1, g.n - 1):       # out-of-bounds starts keep their init exactly
            _, u = g.patch(i)
            assert u == tuple(np.float32(cases[i]))
        for i in (nph * 140 + 10,):     # zero init == the oracle's zero-init run
            _, u = g.patch(i)
            assert np.array_equal(bits(np.array(u)), bits(d["u"][0][i]))
        for i in (nph * 140 + 11, nph * 20 + 5, nph * 159 + 3):
            _, u = g.patch(i)
            assert np.isfinite(u).all()
            # either converged near the zero-init solution or was reset to its init
            near = np.linalg.norm(np.array(u) - d["u"][0][i]) < 0.05
            reset = u == tuple(np.float32(cases[i]))
            assert near or reset, (i, u, d["u"][0][i])
    finally:
        g.close()


def test_grid_state_errors(eng):
    g = Grid(eng, 64, 48, 8)
    try:
        assert g.lib.dis_grid_inverse_search(g.g) == -3          # DIS_ERR_STATE: no images yet
        out = np.empty((48, 64, 2), np.float32)
        assert g.lib.dis_grid_densify(g.g, out.ctypes.data_as(_fp)) == -3
        assert g.lib.dis_grid_set_patch_init(g.g, g.n, C.c_float(0), C.c_float(0)) == -1
        r, u = g.patch(0)
        assert u == (0.0, 0.0)                                     # Patch::reset_patch
    finally:
        g.close()
    bad = C.c_void_p()
    assert eng.lib.dis_grid_create(eng._h, 64, 48, 4, C.byref(bad)) == -1   # img_padding < patch_size


def test_parameter_and_geometry_errors(eng):
    for kw in (dict(patch_size=7), dict(patch_size=2), dict(patch_size=14), dict(coarsest_scale=1, finest_scale=2),
               dict(max_iter=-1), dict(patch_overlap=1.0), dict(patch_overlap=-0.1)):
        with pytest.raises(dis_b200.DisError):
            dis_b200.Engine(0, **kw)
    with pytest.raises(dis_b200.DisError):
        dis_b200.Engine(99)                                        # device ordinal out of range
    with pytest.raises(dis_b200.DisError):
        eng.run_pair(np.zeros((4, 4), np.uint8), np.zeros((4, 4), np.uint8))   # coarsest level < 2x2
    with pytest.raises(dis_b200.DisError):
        eng.run_pyramids(([None] * 4,) * 3, ([None] * 4,) * 3, 8, 64, 48)     # required planes missing
    e2 = dis_b200.Engine(0)
    with pytest.raises(dis_b200.DisError):
        e2.stats()                                                 # DIS_ERR_STATE before any run
    e2.close()


def test_fast_mode_is_within_the_dense_tolerance(oracle_port):
    """conv_eps > 0 (an extension, never the default): dense flow stays within the north-star
    0.01 px mean end-point difference of the exact result while doing a fraction of the work."""
    a, b, gt = dis_b200.synth.make_pair(256, 112, seed=53)
    exact = oracle_port.run_pair_u8(a, b)
    e = dis_b200.Engine(0, conv_eps=1e-4)
    try:
        fast = e.run_pair(a, b)
        st = e.stats()
    finally:
        e.close()
    assert np.linalg.norm(fast - exact, axis=-1).mean() <= TOL_DENSE
    assert st["patch_iterations"] < 0.1 * st["patches"] * 1001
    assert abs(np.linalg.norm(fast - gt, axis=-1).mean() - np.linalg.norm(exact - gt, axis=-1).mean()) < 0.01


def test_flo_roundtrip(tmp_path, eng):
    a, b, _ = dis_b200.synth.make_pair(96, 64, seed=54)
    flow = eng.run_pair(a, b)
    p = str(tmp_path / "x.flo")
    dis_b200.write_flo(p, flow)
    raw = open(p, "rb").read()
    assert raw[:4] == b"PIEH" and np.frombuffer(raw[4:12], np.int32).tolist() == [96, 64]   # src/IO_flow.cpp:60-75
    assert np.array_equal(bits(dis_b200.read_flo(p)), bits(flow))
