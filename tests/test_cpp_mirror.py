"""The C++ mirror of the reference's class API (include/dis_b200.hpp): OpticalFlowClass, PatchGrid, Patch.

tests/cpp/mirror_test.cpp uses the classes the way the reference's own code does
(/root/reference/src/main.cpp:184-189, src/optical_flow.cpp:61-90, src/patch_grid.cpp:68-106); it is
compiled here with plain g++ (no CUDA, Eigen or OpenCV headers) against libdis_b200.so.  Without a GPU
it must fail loudly (no CPU fallback); on the GPU its three routes must reproduce the oracle bit for bit.
"""
import os
import struct
import subprocess

import numpy as np
import pytest

import dis_b200

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "optical-flow-using-dense-inverse-search_b200")


@pytest.fixture(scope="module")
def mirror_exe(tmp_path_factory):
    if not os.path.exists(os.path.join(PKG, "libdis_b200.so")):
        pytest.skip("libdis_b200.so not built")
    exe = str(tmp_path_factory.mktemp("cpp") / "mirror_test")
    subprocess.check_call(["g++", "-std=c++14", "-Wall", "-Wextra", "-Werror", "-O1", "-I", os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tests", "cpp", "mirror_test.cpp"), "-L", PKG, "-ldis_b200",
                           f"-Wl,-rpath,{PKG}", "-o", exe])
    return exe


def write_input(path, pa, pb, W, H, pad, coarsest, finest, iters, ps, patnorm, overlap):
    with open(path, "wb") as f:
        f.write(struct.pack("<8i", W, H, pad, coarsest, finest, iters, ps, int(patnorm)))
        f.write(struct.pack("<f", overlap))
        for pyr in (*pa, *pb):
            for s in range(coarsest + 1):
                f.write(np.ascontiguousarray(pyr[s], np.float32).tobytes())


def test_mirror_fails_loudly_without_a_gpu(mirror_exe, tmp_path, oracle_port):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    a, b, _ = dis_b200.synth.make_pair(64, 48, seed=1)
    fa, _, _ = oracle_port.pad_convert(a, 1)
    fb, _, _ = oracle_port.pad_convert(b, 1)
    H, W = fa.shape
    write_input(tmp_path / "in.bin", oracle_port.pyramid(fa, 2, 8), oracle_port.pyramid(fb, 2, 8), W, H, 8, 1, 0, 5, 8, True, 0.7)
    r = subprocess.run([mirror_exe, str(tmp_path / "in.bin"), str(tmp_path / "out.bin")], capture_output=True, text=True)
    assert r.returncode == 3 and "no CUDA device" in r.stderr
    assert not os.path.exists(tmp_path / "out.bin")


@pytest.mark.gpu
@pytest.mark.parametrize("ps,overlap,coarsest,finest,iters", [(8, 0.7, 3, 0, 1000), (12, 0.75, 2, 0, 40), (8, 0.7, 2, 1, 30)])
def test_mirror_classes_reproduce_the_oracle(mirror_exe, tmp_path, oracle_port, ps, overlap, coarsest, finest, iters):
    a, b, _ = dis_b200.synth.make_pair(256, 112, seed=21)
    fa, _, _ = oracle_port.pad_convert(a, coarsest)
    fb, _, _ = oracle_port.pad_convert(b, coarsest)
    H, W = fa.shape
    pa, pb = oracle_port.pyramid(fa, coarsest + 1, ps), oracle_port.pyramid(fb, coarsest + 1, ps)
    ref, d = oracle_port.optical_flow(pa, pb, ps, W, H, coarsest=coarsest, finest=finest, iterations=iters, patch_size=ps,
                                      overlap=overlap, dump=True)
    write_input(tmp_path / "in.bin", pa, pb, W, H, ps, coarsest, finest, iters, ps, True, overlap)
    r = subprocess.run([mirror_exe, str(tmp_path / "in.bin"), str(tmp_path / "out.bin")], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    # the constructor prints one "scale: N" line per level, coarse to fine (src/optical_flow.cpp:69)
    assert r.stdout.split() == [t for s in range(coarsest, finest - 1, -1) for t in ("scale:", str(s))]
    raw = open(tmp_path / "out.bin", "rb").read()
    n = ref.size
    ctor = np.frombuffer(raw, np.float32, n).reshape(ref.shape)
    grid = np.frombuffer(raw, np.float32, n, 4 * n).reshape(ref.shape)
    assert np.array_equal(ctor.view(np.uint32), ref.view(np.uint32)), "OpticalFlowClass differs from the oracle"
    assert np.array_equal(grid.view(np.uint32), ref.view(np.uint32)), "PatchGrid level loop differs from the oracle"
    (k,) = struct.unpack_from("<i", raw, 8 * n)
    assert k >= 10
    u_top = d["u"][coarsest - finest]
    for i in range(k):
        pid, ux, uy = struct.unpack_from("<iff", raw, 8 * n + 4 + 12 * i)
        assert np.array_equal(np.float32([ux, uy]).view(np.uint32), u_top[pid].view(np.uint32)), f"Patch {pid} differs"
