"""dis_pool_*: several GPUs of one box behind one handle — the pair loop of /root/reference/src/main.cpp:102 sharded over
devices (contiguous blocks, one shared frame between neighbouring blocks of a sequence, no collective).  The sharding
rule is checked on the CPU; the GPU tests run two engines on ONE GPU (devices [0, 0]) so that the pool is exercised on a
single-GPU box too, and two different GPUs where the box has them."""
import ctypes as C

import numpy as np
import pytest

import dis_b200


def test_pool_shard_rule_matches_the_python_plumbing():
    lib = dis_b200.load_library()
    lib.dis_pool_shard.restype = None
    for n in (0, 1, 2, 7, 49, 50, 4096):
        for parts in (1, 2, 3, 4, 8):
            prev = 0
            for r in range(parts):
                lo, hi = C.c_int(), C.c_int()
                lib.dis_pool_shard(n, parts, r, C.byref(lo), C.byref(hi))
                assert (lo.value, hi.value) == dis_b200.parallel.shard_pairs(n, parts, r)
                assert lo.value == prev
                prev = hi.value
            assert prev == n


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


@pytest.fixture(scope="module")
def seq():
    frames, _ = dis_b200.synth.make_sequence(12, 200, 100, seed=21)
    e = dis_b200.Engine(0)
    ref = e.run_sequence(frames)
    st = e.stats()
    e.close()
    return frames, ref, st


def devices_for_test():
    import torch
    out = [[0, 0], [0, 0, 0]]
    if torch.cuda.device_count() >= 2:
        out.append([0, 1])
    if torch.cuda.device_count() >= 4:
        out.append([0, 1, 2, 3])
    return out


@pytest.mark.gpu
def test_pool_sequence_and_batch_equal_one_engine(seq):
    frames, ref, st = seq
    for devs in devices_for_test():
        p = dis_b200.Pool(devs)
        try:
            out = p.run_sequence(frames)
            assert np.array_equal(bits(out), bits(ref)), f"sequence over devices {devs}"
            ps = p.stats()
            assert ps["patches"] == st["patches"] and ps["patch_iterations"] == st["patch_iterations"]
            wall, per = p.last_ms()
            assert wall > 0 and len(per) == len(devs) and all(x > 0 for x in per)
            outb = p.run_batch(frames[:-1], frames[1:])
            assert np.array_equal(bits(outb), bits(ref)), f"batch over devices {devs}"
            assert [w["device"] for w in p.workers()] == devs
            # fewer pairs than workers: some workers get nothing
            few = p.run_sequence(frames[:2])
            assert np.array_equal(bits(few[0]), bits(ref[0]))
        finally:
            p.close()


@pytest.mark.gpu
def test_pool_pinned_buffers_and_errors(seq):
    frames, ref, _ = seq
    n, h, w = frames.shape
    p = dis_b200.Pool([0, 0])
    try:
        src = p.host_alloc(0, frames.nbytes)
        dst = p.host_alloc(1, ref.nbytes)
        C.memmove(src, frames.ctypes.data, frames.nbytes)
        p.run_sequence_ptr(src, n, w, h, dst)
        got = np.ctypeslib.as_array(C.cast(dst, C.POINTER(C.c_float)), shape=ref.shape)
        assert np.array_equal(bits(got), bits(ref))
        p.host_free(src); p.host_free(dst)
        with pytest.raises(dis_b200.DisError):
            p.run_sequence(frames[:1])
    finally:
        p.close()
    with pytest.raises(dis_b200.DisError):
        dis_b200.Pool([99])


@pytest.mark.gpu
def test_pool_async_submissions_equal_blocking_calls(seq):
    """dis_pool_submit_* / dis_pool_wait: whole calls queued to the engines (two engines on one GPU hide each other's
    copies); flows, stats and errors are those of the blocking call, in any waiting order."""
    frames, ref, st = seq
    n, h, w = frames.shape
    for devs in devices_for_test():
        p = dis_b200.Pool(devs)
        try:
            outs = [np.zeros_like(ref) for _ in range(5)]
            tickets = [p.submit_sequence(frames, o) for o in outs[:4]]
            tickets.append(p.submit_batch(frames[:-1], frames[1:], outs[4]))
            assert len(set(tickets)) == 5
            workers = []
            for t, o in reversed(list(zip(tickets, outs))):            # waited for in reverse order
                s = p.wait(t)
                assert np.array_equal(bits(o), bits(ref)), f"ticket {t} over devices {devs}"
                assert s["patches"] == st["patches"] and s["patch_iterations"] == st["patch_iterations"]
                workers.append(s["worker"])
            assert len(set(workers)) == min(len(devs), 5)              # equal jobs spread over all the engines
            with pytest.raises(dis_b200.DisError):
                p.wait(tickets[0])                                     # a ticket is waited for once
            with pytest.raises(dis_b200.DisError):
                p.wait(12345)
            with pytest.raises(dis_b200.DisError):
                p.submit_sequence(frames[:1], np.zeros((0, h, w, 2), np.float32))
            # a failing job reports through its ticket, the pool stays usable; blocking calls queue behind submissions
            big = dis_b200.Pool(devs, coarsest_scale=9)
            try:
                small = np.zeros((2, 64, 64), np.uint8)
                t = big.submit_sequence(small, np.zeros((1, 64, 64, 2), np.float32))
                with pytest.raises(dis_b200.DisError):
                    big.wait(t)
            finally:
                big.close()
            t1 = p.submit_sequence(frames, outs[0])
            again = p.run_sequence(frames)
            assert np.array_equal(bits(again), bits(ref))
            t2 = p.submit_sequence(frames, outs[1])
            p.wait_all()
            assert np.array_equal(bits(outs[0]), bits(ref)) and np.array_equal(bits(outs[1]), bits(ref))
            with pytest.raises(dis_b200.DisError):
                p.wait(t2)
            del t1
        finally:
            p.close()


@pytest.mark.gpu
def test_host_slices_switch_does_not_change_results():
    frames, _ = dis_b200.synth.make_sequence(10, 200, 100, seed=5)
    e = dis_b200.Engine(0)
    a = e.run_sequence(frames)
    e.set_host_slices(False)
    b = e.run_sequence(frames)
    e.set_host_slices(True)
    c = e.run_sequence(frames)
    e.close()
    assert np.array_equal(bits(a), bits(b)) and np.array_equal(bits(a), bits(c))
