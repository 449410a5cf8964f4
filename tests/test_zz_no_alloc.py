"""dis_color_code_device_noalloc (include/dis_b200.h): the colour coding for callers that cannot allocate.

No counterpart in the reference; the check is against the existing
entry point dis_color_code, byte for byte (integer output: the tolerance of the colour tests is 0), on two calls in a
row that share the library's fixed scratch.  Named test_zz_* so that pytest collects it last.
"""
import ctypes as C

import numpy as np
import pytest

import dis_b200
from dis_b200 import engine as E


def _fields():
    rng = np.random.default_rng(11)
    big = (rng.normal(size=(301, 517, 2)) * 9).astype(np.float32)        # 38 partials, max |u| ~ 40
    big[0, 0] = np.nan; big[5, 7] = (3e9, 1); big[9, 9] = 0; big[10, 10] = (0, -0.0)
    small = (rng.normal(size=(97, 131, 2)) * 0.8).astype(np.float32)      # 4 partials, max |u| ~ 3: stale partials of
    return big, small                                                    # the first call must not leak into the second


@pytest.mark.gpu
def test_noalloc_colour_matches_existing_entry_point_twice():
    import torch
    assert torch.cuda.is_available()
    dev = 0
    eng = dis_b200.Engine(device=dev)
    big, small = _fields()
    # the caller owns inputs and outputs (allocated BEFORE the calls under test)
    d_in = [torch.from_numpy(f).to(f"cuda:{dev}") for f in (big, small)]
    d_out = [torch.full(f.shape[:2] + (3,), 77, dtype=torch.uint8, device=f"cuda:{dev}") for f in (big, small)]
    torch.cuda.synchronize(dev)
    stream = torch.cuda.Stream(device=dev)
    for rep in range(2):                                                  # and again: contents of the last call stay behind
        for f, di, do in zip((big, small), d_in, d_out):
            h, w, _ = f.shape
            E.color_code_device_noalloc(dev, di.data_ptr(), w, h, do.data_ptr(), -1.0, stream.cuda_stream)
        stream.synchronize()
        for f, do in zip((big, small), d_out):
            assert np.array_equal(do.cpu().numpy(), eng.color_code(f)), f"rep {rep}, field {f.shape}"
            do.fill_(77)
        torch.cuda.synchronize(dev)
    # fixed range: the reduction is skipped, scratch untouched
    h, w, _ = small.shape
    E.color_code_device_noalloc(dev, d_in[1].data_ptr(), w, h, d_out[1].data_ptr(), 1.5, stream.cuda_stream)
    stream.synchronize()
    assert np.array_equal(d_out[1].cpu().numpy(), eng.color_code(small, 1.5))
    # default stream (NULL)
    E.color_code_device_noalloc(dev, d_in[0].data_ptr(), big.shape[1], big.shape[0], d_out[0].data_ptr())
    torch.cuda.synchronize(dev)
    assert np.array_equal(d_out[0].cpu().numpy(), eng.color_code(big))
    eng.close()


@pytest.mark.gpu
def test_noalloc_colour_allocates_nothing():
    import torch
    dev = 0
    f = _fields()[1]
    h, w, _ = f.shape
    di = torch.from_numpy(f).to(f"cuda:{dev}")
    do = torch.empty((h, w, 3), dtype=torch.uint8, device=f"cuda:{dev}")
    E.color_code_device_noalloc(dev, di.data_ptr(), w, h, do.data_ptr())   # set-up (symbol address, module load) done
    torch.cuda.synchronize(dev)
    free0, _ = torch.cuda.mem_get_info(dev)
    for _ in range(3):
        E.color_code_device_noalloc(dev, di.data_ptr(), w, h, do.data_ptr())
    torch.cuda.synchronize(dev)
    free1, _ = torch.cuda.mem_get_info(dev)
    assert free1 >= free0          # nothing was taken from the device between the two readings


def test_noalloc_colour_rejects_bad_sizes():
    """Argument checks come before any CUDA call: they run without a GPU."""
    lib = E.load_library()
    call = lib.dis_color_code_device_noalloc
    one = C.c_void_p(256)   # never dereferenced: the size check fails first
    assert E.NOALLOC_MAX_PIXELS == 8192 * 4096
    rc = call(C.c_int(0), one, C.c_int(8192), C.c_int(4097), C.c_float(-1.0), one, C.c_void_p(0))
    assert rc == -1
    msg = lib.dis_last_error().decode()
    assert "exceeds the fixed scratch" in msg and "33554432" in msg
    assert call(C.c_int(0), one, C.c_int(0), C.c_int(4), C.c_float(-1.0), one, C.c_void_p(0)) == -1
    assert call(C.c_int(0), C.c_void_p(0), C.c_int(4), C.c_int(4), C.c_float(-1.0), one, C.c_void_p(0)) == -1
    assert call(C.c_int(99), one, C.c_int(4), C.c_int(4), C.c_float(-1.0), one, C.c_void_p(0)) == -1
