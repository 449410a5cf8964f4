"""Host-side multi-GPU logic on CPU: sharding rule and the gloo world-size-2 statistics reduction
(the GPU job uses the same code over NCCL; no pixel data crosses ranks — SURVEY.md §8e)."""
import os
import socket
import sys

import pytest
import torch.multiprocessing as mp

import dis_b200
from dis_b200.parallel import shard_frames, shard_pairs

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("n,world", [(49, 1), (49, 2), (49, 4), (49, 8), (4096, 8), (3, 8), (0, 2)])
def test_shards_partition_the_pairs(n, world):
    blocks = [shard_pairs(n, world, r) for r in range(world)]
    assert blocks[0][0] == 0 and blocks[-1][1] == n
    assert all(blocks[i][1] == blocks[i + 1][0] for i in range(world - 1))
    sizes = [hi - lo for lo, hi in blocks]
    assert max(sizes) - min(sizes) <= 1 and sizes == sorted(sizes, reverse=True)


def test_sequence_shards_overlap_by_one_frame():
    fr = [shard_frames(50, 4, r) for r in range(4)]
    assert fr == [(0, 14), (13, 26), (25, 38), (37, 50)]
    assert shard_frames(2, 4, 3) == (0, 0)
    with pytest.raises(ValueError):
        shard_pairs(4, 2, 2)


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from dis_b200.parallel import gather_results, reduce_stats
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = shard_pairs(49, world, rank)
    st = {"patches": (hi - lo) * 149632, "patch_iterations": (hi - lo) * 1000, "sampled_iterations": hi - lo,
          "resets": rank, "start_oob": 0, "kernel_launches": 25, "sampled_converge": 0, "sampled_walk": 0}
    red = reduce_stats(st, 10.0 + rank, dist)
    got = gather_results(list(range(lo, hi)), dist)
    dist.barrier()
    dist.destroy_process_group()
    q.put((rank, red, got))


def test_gloo_world2_reduction_and_gather():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    res = sorted(q.get(timeout=120) for _ in ps)
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, red, got in res:
        assert red["patches"] == 49 * 149632 and red["resets"] == 1 and red["kernel_launches"] == 50
        assert red["device_ms"] == 11.0                       # MAX over ranks
        assert [x for part in got for x in part] == list(range(49))
    assert dis_b200.parallel.reduce_stats({k: 1 for k in dis_b200.parallel.STAT_KEYS}, 2.5)["device_ms"] == 2.5


def test_band_rows_partition_the_grid():
    from dis_b200.parallel import band_rows
    for nph, world in [(1080, 8), (28, 8), (3, 8), (220, 2)]:
        bands = [band_rows(nph, world, r) for r in range(world)]
        assert bands[0][0] == 0 and bands[-1][1] == nph
        assert all(bands[i][1] == bands[i + 1][0] for i in range(world - 1))
