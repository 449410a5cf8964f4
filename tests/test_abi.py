"""CPU tests of the drop-in boundary: libdis_b200.so loads, exports exactly the entry points
include/dis_b200.h declares, and — with no GPU — fails loudly instead of falling back to a CPU path."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

import dis_b200
from dis_b200 import engine

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "dis_b200.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"DIS_API\s+[\w\s\*]+?\b(dis_\w+)\s*\(", src)))


def test_header_and_binding_list_agree():
    assert declared_symbols() == sorted(engine.ABI_SYMBOLS)


def test_library_exports_every_declared_symbol_and_nothing_else():
    lib = dis_b200.load_library()
    for s in declared_symbols():
        assert hasattr(lib, s), s
    out = subprocess.run(["nm", "-D", "--defined-only", engine.LIB_PATH], capture_output=True, text=True).stdout
    exported = sorted(l.split()[-1] for l in out.splitlines() if " T " in l)
    assert [s for s in exported if s.startswith("dis_")] == declared_symbols()
    # -fvisibility=hidden: no kernel launchers or helpers leak out of the C ABI
    assert not [s for s in exported if not s.startswith("dis_") and not s.startswith("_")], exported


def test_default_params_are_main_cpp_defaults():
    p = dis_b200.default_params()      # src/main.cpp:63-72
    assert (p.max_iter, p.patch_size, p.coarsest_scale, p.finest_scale, p.patch_norm) == (1000, 8, 3, 0, 1)
    assert p.patch_overlap == np.float32(0.7) and p.conv_eps == 0.0
    assert C.sizeof(engine.DisParams) == 48 and C.sizeof(engine.DisLevelInfo) == 44 and C.sizeof(engine.DisStats) == 64
    assert b"sm_100a" in dis_b200.load_library().dis_version()


def _no_gpu():
    import torch
    return not torch.cuda.is_available()


@pytest.mark.skipif(not _no_gpu(), reason="checks the behaviour of a box WITHOUT a GPU")
def test_no_gpu_means_error_not_fallback():
    with pytest.raises(dis_b200.DisError, match="no CUDA device|CUDA"):
        dis_b200.Engine(0)
    lib = dis_b200.load_library()
    h = C.c_void_p()
    p = dis_b200.default_params()
    assert lib.dis_create(C.byref(p), 0, C.byref(h)) == -2     # DIS_ERR_CUDA
    assert not h.value
    assert b"no CPU path" in lib.dis_last_error()


def test_parameter_validation_precedes_device_use():
    lib = dis_b200.load_library()
    h = C.c_void_p()
    for kw in (dict(patch_size=7), dict(patch_size=18), dict(coarsest_scale=0, finest_scale=1), dict(max_iter=-5),
               dict(patch_overlap=1.5)):
        p = dis_b200.default_params(**kw)
        assert lib.dis_create(C.byref(p), 0, C.byref(h)) == -1, kw     # DIS_ERR_PARAM
    assert lib.dis_create(None, 0, C.byref(h)) == -1
    assert lib.dis_num_levels(None) == 0
    assert lib.dis_grid_num_patches(None) == 0


def test_flo_io_is_host_only(tmp_path):
    """Middlebury .flo (src/IO_flow.cpp:10-98): 'PIEH', int32 w, h, interleaved fp32 rows."""
    rng = np.random.default_rng(1)
    flow = rng.standard_normal((7, 13, 2)).astype(np.float32)
    p = str(tmp_path / "a.flo")
    dis_b200.write_flo(p, flow)
    raw = open(p, "rb").read()
    assert len(raw) == 12 + flow.nbytes and raw[:4] == b"PIEH"
    assert np.frombuffer(raw[:4], np.float32)[0] == np.float32(202021.25)    # TAG_FLOAT
    assert np.array_equal(dis_b200.read_flo(p), flow)
    open(p, "wb").write(b"XXXX" + raw[4:])
    with pytest.raises(dis_b200.DisError):
        dis_b200.read_flo(p)
    with pytest.raises(dis_b200.DisError):
        dis_b200.read_flo(str(tmp_path / "missing.flo"))


def test_product_never_touches_the_oracle():
    pkg = os.path.join(ROOT, "optical-flow-using-dense-inverse-search_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".hpp")):
                assert "oracle" not in open(os.path.join(dp, f), errors="ignore").read().lower(), f


def test_packed_arithmetic_is_not_contracted():
    """ptxas 12.9 contracts mul.rn.f32x2 + add.rn.f32x2 into one FFMA2 (it honours the rounding modifiers and
    --fmad=false only for scalar code), which would break bit parity.  csrc/dis_search.cu therefore writes every packed
    addition as fma(a, ONE, b) with a run-time ONE.  Check the SASS of the build: a sample of k_converge<8,4> must be
    exactly 48 packed multiplications (4 x 8 bilinear + 2 x 8 projections) and 53 packed additions (3 x 8 bilinear,
    8 mean subtractions, 3 x 7 tree sums) — a contracted pair would show up as one FFMA2 less and one FMUL2 less."""
    import shutil
    import subprocess
    if shutil.which("cuobjdump") is None:
        pytest.skip("cuobjdump not available")
    lib = os.path.join(ROOT, "optical-flow-using-dense-inverse-search_b200", "libdis_b200.so")
    names = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
    fn, body = None, []
    for line in names.splitlines():
        if "Function :" in line:
            fn = line.split("Function :")[1].strip()
        elif fn and "k_convergeILi8ELi4ELb0" in fn:
            body.append(line)
    assert body, "k_converge<8, 4, false> not found in the library"
    text = "\n".join(body)
    n_mul = sum(1 for l in body if " FMUL2 " in l)
    n_fma = sum(1 for l in body if " FFMA2 " in l)
    # (the compiler keeps an out-of-line copy of one 7-addition tree sum for a warp that reaches the shuffles diverged:
    #  60 FFMA2 in all; a contracted pair would remove one FMUL2)
    assert n_mul == 48 and n_fma >= 53, f"packed arithmetic contracted or restructured: FMUL2 {n_mul}, FFMA2 {n_fma}"
    assert " FADD2 " not in text            # every packed addition goes through the opaque-one FFMA2
