"""One large frame tiled by grid rows over several GPUs (BASELINE.json config 5): the band
kernels, the per-level exchange of patch flows and the redundant densification must reproduce the
single-engine run bit for bit.  The N ranks are emulated by N engines on cuda:0 (the driver's GPU tier has
one GPU); with >= 2 GPUs the same schedule also runs over NCCL with one process per GPU."""
import os
import socket
import sys

import numpy as np
import pytest

import dis_b200
from dis_b200.parallel import band_rows, run_pair_tiled_emulated

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


@pytest.mark.parametrize("world,w,h,kw", [
    (2, 512, 256, {}),
    (3, 250, 107, dict(max_iter=40)),                                   # ragged size, bands of unequal height
    (4, 480, 272, dict(patch_size=12, patch_overlap=0.75, coarsest_scale=2, max_iter=60)),
    (8, 128, 64, dict(max_iter=30)),                                    # more ranks than grid rows at the coarsest level
])
def test_tiled_emulated_matches_single_engine(world, w, h, kw):
    a, b, _ = dis_b200.synth.make_pair(w, h, seed=31, amp=4.0)
    single = dis_b200.Engine(0, **kw)
    engines = [dis_b200.Engine(0, **kw) for _ in range(world)]
    try:
        ref = single.run_pair(a, b)
        st_ref = single.stats()
        flows = run_pair_tiled_emulated(engines, a, b)
        for f in flows:
            assert np.array_equal(bits(f), bits(ref))
        # the bands partition the work: iteration counters add up to the single-engine run
        assert sum(e.stats()["patch_iterations"] for e in engines) == st_ref["patch_iterations"]
        nph = engines[0].level_info(0)["nph"]
        assert [band_rows(nph, world, r) for r in range(world)][-1][1] == nph
    finally:
        single.close()
        for e in engines:
            e.close()


def test_tiled_4k_bands_match_single_engine():
    """config 5 shape: 3840x2160, two bands, a short iteration budget that crosses the walk hand-off."""
    a, b, _ = dis_b200.synth.make_pair(3840, 2160, seed=2, amp=11.0)
    single = dis_b200.Engine(0, max_iter=30)
    engines = [dis_b200.Engine(0, max_iter=30) for _ in range(2)]
    try:
        ref = single.run_pair(a, b)
        for f in run_pair_tiled_emulated(engines, a, b):
            assert np.array_equal(bits(f), bits(ref))
    finally:
        single.close()
        for e in engines:
            e.close()


def test_tiled_call_order_is_checked():
    e = dis_b200.Engine(0)
    try:
        with pytest.raises(dis_b200.DisError):
            e.tiled_search(0, 0, 1)
        a, b, _ = dis_b200.synth.make_pair(128, 64, seed=1)
        e.tiled_begin(a, b)
        with pytest.raises(dis_b200.DisError):
            e.tiled_search(0, 0, 10 ** 6)
        with pytest.raises(dis_b200.DisError):
            e.tiled_search(99, 0, 1)
    finally:
        e.close()


def _nccl_worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    import dis_b200 as d
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    a, b, _ = d.synth.make_pair(1024, 436, seed=5)
    e = d.Engine(rank, max_iter=100)
    flow = d.parallel.run_pair_tiled(e, a, b, rank, world, dist)
    ref = e.run_pair(a, b)
    e.close()
    dist.barrier()
    dist.destroy_process_group()
    q.put((rank, bool(np.array_equal(flow.view(np.uint32), ref.view(np.uint32)))))


def test_tiled_over_nccl_when_two_gpus_are_present():
    import torch
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run with gpurun --gpus 2)")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_nccl_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = sorted(q.get(timeout=300) for _ in ps)
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res == [(0, True), (1, True)]


# ---- the in-process path: dis_pool_run_pair_tiled_u8 — bands of grid rows, HALO exchange over peer memory, every GPU
#      densifies and downloads only its own band (csrc/dis_pool.cu) ------------------------------------------------------
def pool_device_lists(n):
    import torch
    out = [[0] * n]                                   # n engines on one GPU: the exchange degenerates to local copies
    if torch.cuda.device_count() >= n > 1:
        out.append(list(range(n)))                    # real peers over NVLink
    return out


@pytest.mark.parametrize("world,w,h,kw", [
    (1, 256, 112, {}),
    (2, 512, 256, {}),
    (3, 250, 107, dict(max_iter=40)),                                   # ragged size, bands of unequal height
    (4, 480, 272, dict(patch_size=12, patch_overlap=0.75, coarsest_scale=2, max_iter=60)),
    (8, 128, 64, dict(max_iter=30)),                                    # more GPUs than grid rows at the coarsest level
    (5, 200, 100, dict(patch_size=4, max_iter=25)),
])
def test_pool_tiled_halo_exchange_matches_single_engine(world, w, h, kw):
    a, b, _ = dis_b200.synth.make_pair(w, h, seed=33, amp=4.0)
    single = dis_b200.Engine(0, **kw)
    try:
        ref = single.run_pair(a, b)
        st_ref = single.stats()
    finally:
        single.close()
    for devs in pool_device_lists(world):
        p = dis_b200.Pool(devs, **kw)
        try:
            out = p.run_pair_tiled(a, b)
            assert np.array_equal(bits(out), bits(ref)), f"devices {devs}"
            again = p.run_pair_tiled(a, b)
            assert np.array_equal(bits(again), bits(ref)), f"second run, devices {devs}"
            st = p.stats()
            # the bands partition the patches and the iteration counters add up to the single-engine run
            assert st["patches"] == st_ref["patches"] and st["patch_iterations"] == st_ref["patch_iterations"]
        finally:
            p.close()


def test_pool_tiled_4k_matches_single_engine():
    """config 5 shape, three bands, a budget that crosses the walk hand-off."""
    a, b, _ = dis_b200.synth.make_pair(3840, 2160, seed=2, amp=11.0)
    single = dis_b200.Engine(0, max_iter=30)
    try:
        ref = single.run_pair(a, b)
    finally:
        single.close()
    for devs in pool_device_lists(3):
        p = dis_b200.Pool(devs, max_iter=30)
        try:
            assert np.array_equal(bits(p.run_pair_tiled(a, b)), bits(ref)), f"devices {devs}"
        finally:
            p.close()


def test_pool_tiled_rejects_finest_scale():
    a, b, _ = dis_b200.synth.make_pair(128, 64, seed=1)
    p = dis_b200.Pool([0, 0], finest_scale=1)
    try:
        with pytest.raises(dis_b200.DisError):
            p.run_pair_tiled(a, b)
    finally:
        p.close()
