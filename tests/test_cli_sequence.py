"""tools/dis_sequence.cpp — the reference's command line (/root/reference/src/main.cpp:60-111) as a
headless sequence driver over the C ABI: argument forms, file naming, .flo and colour-coded outputs."""
import os
import subprocess

import numpy as np
import pytest

import dis_b200

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "optical-flow-using-dense-inverse-search_b200")


@pytest.fixture(scope="module")
def cli(tmp_path_factory):
    if not os.path.exists(os.path.join(PKG, "libdis_b200.so")):
        pytest.skip("libdis_b200.so not built")
    exe = str(tmp_path_factory.mktemp("cli") / "dis_sequence")
    subprocess.check_call(["g++", "-std=c++14", "-Wall", "-Wextra", "-Werror", "-O1", "-I", os.path.join(ROOT, "include"),
                           os.path.join(ROOT, "tools", "dis_sequence.cpp"), "-L", PKG, "-ldis_b200", f"-Wl,-rpath,{PKG}", "-o", exe])
    return exe


def write_pgm(path, img):
    with open(path, "wb") as f:
        f.write(b"P5\n# synthetic\n%d %d\n255\n" % (img.shape[1], img.shape[0]))
        f.write(np.ascontiguousarray(img, np.uint8).tobytes())


def test_cli_argument_forms(cli, tmp_path):
    r = subprocess.run([cli, "a", "b"], capture_output=True, text=True, cwd=tmp_path)
    assert r.returncode == 0 and r.stdout.startswith("Not good parameters!")          # main.cpp:93-101
    r = subprocess.run([cli, "nowhere", "1", "3"], capture_output=True, text=True, cwd=tmp_path)
    assert r.returncode == 255 and "No image data" in r.stdout                        # main.cpp:118-122 (return -1)
    r = subprocess.run([cli, "nowhere", "5", "5"], capture_output=True, text=True, cwd=tmp_path)
    assert r.returncode == 0 and r.stdout == ""                                       # empty frame range: the loop never runs


@pytest.mark.gpu
def test_cli_sequence_outputs(cli, tmp_path, oracle_port):
    frames, _ = dis_b200.synth.make_sequence(4, 128, 64, seed=11)
    os.makedirs(tmp_path / "seq")
    for i, fr in enumerate(frames):
        write_pgm(tmp_path / "seq" / f"frame_{i + 3:04d}.pgm", fr)
    args = ["seq", "3", "6", "60", "8", "2", "0", "0.7", "1", "0"]
    r = subprocess.run([cli] + args, capture_output=True, text=True, cwd=tmp_path)
    assert r.returncode == 0, r.stderr
    assert r.stdout.count("start seq/frame_") == 3 and r.stdout.count("finish seq/frame_") == 3
    e = dis_b200.Engine(0, max_iter=60, coarsest_scale=2)
    try:
        ref = e.run_sequence(frames)
    finally:
        e.close()
    for i in range(3):
        flo = dis_b200.read_flo(str(tmp_path / "OF_seq" / f"frame_{i + 3:04d}.flo"))
        assert np.array_equal(flo.view(np.uint32), ref[i].view(np.uint32))
        raw = open(tmp_path / "OF_seq" / f"frame_{i + 3:04d}.ppm", "rb").read()
        assert raw.startswith(b"P6\n128 64\n255\n")
        rgb = np.frombuffer(raw[-3 * 128 * 64:], np.uint8).reshape(64, 128, 3)
        assert np.array_equal(rgb[..., ::-1], oracle_port.color_code(ref[i]))


@pytest.mark.gpu
def test_cli_sequence_longer_than_one_chunk(cli, tmp_path, oracle_port):
    """131 frames = 130 pairs: the engine works in chunks of 128 pairs; every output must still belong to its own pair."""
    frames, _ = dis_b200.synth.make_sequence(131, 64, 48, seed=12)
    os.makedirs(tmp_path / "long")
    for i, fr in enumerate(frames):
        write_pgm(tmp_path / "long" / f"frame_{i + 1:04d}.pgm", fr)
    r = subprocess.run([cli, "long", "1", "131", "30", "8", "2", "0", "0.7", "1", "0"], capture_output=True, text=True, cwd=tmp_path)
    assert r.returncode == 0, r.stderr
    assert r.stdout.count("finish long/frame_") == 130
    e = dis_b200.Engine(0, max_iter=30, coarsest_scale=2)
    try:
        for i in (0, 1, 127, 128, 129):
            ref = e.run_pair(frames[i], frames[i + 1])
            flo = dis_b200.read_flo(str(tmp_path / "OF_long" / f"frame_{i + 1:04d}.flo"))
            assert np.array_equal(flo.view(np.uint32), ref.view(np.uint32)), f"pair {i}"
            raw = open(tmp_path / "OF_long" / f"frame_{i + 1:04d}.ppm", "rb").read()
            rgb = np.frombuffer(raw[-3 * 64 * 48:], np.uint8).reshape(48, 64, 3)
            assert np.array_equal(rgb[..., ::-1], oracle_port.color_code(ref)), f"colour image of pair {i}"
    finally:
        e.close()
