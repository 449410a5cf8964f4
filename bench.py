#!/usr/bin/env python
"""bench.py — frame pairs/s at Sintel shape (BASELINE.json metric) for the B200 DIS engine.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

A step = one pass of the whole hot path (pyramids -> 4 levels of inverse search + densification ->
crop) over the workload of BASELINE.json configs[1]: a 50-frame synthetic 1024x436 sequence = 49
frame pairs, default parameters (patch 8, scales 3->0, overlap 0.7, mean-norm on, max_iter 1000),
EXACT mode (every one of the reference's 1001 passes per patch is accounted for, bit-identically).
N > 1: one process per GPU (torchrun), each rank runs its own sequence (weak scaling, no collective
on the data path; NCCL only for the barrier and the max-over-ranks of the device time).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

W, H, N_FRAMES = 1024, 436, 50
N_PAIRS = N_FRAMES - 1
METRIC = "frame pairs/s @1024x436"
UNIT = "pairs/s"
PS = 8
# one patch-iteration = 16*ps^2 tap bytes and 13*ps^2+32 fp32 ops
B_TAP = 16 * PS * PS
F_ITER = 13 * PS * PS + 32


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), float(d.get("sm_max_mhz", 1965.0)), "measured"
    return 6650.0, 1965.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.proc = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            pass

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            out, _ = self.proc.communicate(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill(); out, _ = self.proc.communicate()
        sm, mx, reasons, pw = [], [], set(), []
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for line in out.strip().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


def strip_patches(oracle_obj, w, h):
    """patches summed over the 4 levels for a w x h frame (work is proportional to this)."""
    Wp, Hp = w + oracle_obj.div_pad(w, 3), h + oracle_obj.div_pad(h, 3)
    return sum((lambda i: i["npw"] * i["nph"])(oracle_obj.level_info(Wp, Hp, s, PS, PS, 0.7)) for s in range(4))


_SEQ = {}


def bench_sequence(seed: int):
    """The timed workload: make_sequence(50, 1024, 436, seed) — cached, both arms and the parity check use it."""
    import dis_b200
    if seed not in _SEQ:
        _SEQ[seed] = dis_b200.synth.make_sequence(N_FRAMES, W, H, seed=seed)
    return _SEQ[seed]


def cpu_run(kind: str, threads: int, strip_h: int, first_pair: int):
    """Time `threads` concurrent single-threaded runs of the CPU reference (one pair each, as the reference is
    single-threaded) on pairs first_pair, first_pair + 1, ... (mod 49) of the timed sequence, optionally cropped
    to their top strip_h rows.  Returns (equivalent full pairs/s, secs, fraction of a pair one run is)."""
    import oracle
    orc = oracle.Oracle(kind)
    frames, _ = bench_sequence(100)
    pairs = []
    for t in range(threads):
        i = (first_pair + t) % N_PAIRS
        pairs.append((np.ascontiguousarray(frames[i][:strip_h]), np.ascontiguousarray(frames[i + 1][:strip_h])))
    equiv = 1.0 if strip_h >= H else strip_patches(orc, W, strip_h) / strip_patches(orc, W, H)
    ths = [threading.Thread(target=orc.run_pair_u8, args=p) for p in pairs]
    t0 = time.perf_counter()
    for t in ths: t.start()
    for t in ths: t.join()
    dt = time.perf_counter() - t0
    return threads * equiv / dt, dt, equiv


def reference_arm(args, rank: int):
    """--impl reference: the reference's own CPU implementation of the path on the host cores (oracle/_ref = its
    unmodified engine sources; else the C port), rank 0 only.  Same config as the GPU arm: WHOLE 1024x436 pairs of
    the same 50-frame sequence, default parameters, max_iter 1000; a step = one pair per host core, all cores busy.
    CPU code needs no warm-up, so the --warmup steps are not run (recorded in the line); --ref-strip H crops the
    pairs to their top H rows for a quick look (recorded as same_config: false)."""
    if rank != 0:
        return
    import oracle
    kind = "reference" if oracle.load_ref() is not None else "port"
    cores = os.cpu_count() or 1
    strip_h = args.ref_strip if 0 < args.ref_strip < H else H
    secs, equiv = [], 1.0
    for s in range(args.steps):
        _, dt, equiv = cpu_run(kind, cores, strip_h, first_pair=s * cores)
        secs.append(dt)
    t = float(np.sum(secs))
    value = cores * equiv * len(secs) / t
    whole = strip_h >= H
    sample = (f"each step: {cores} concurrent single-threaded runs, one pair of the timed 50-frame sequence each, "
              + ("whole 1024x436 pairs" if whole else f"cropped to 1024x{strip_h} (= {equiv:.4f} of a pair by patch count)")
              + ", default parameters, max_iter 1000; no warm-up steps")
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": 0, "warmup_requested": args.warmup, "ms_per_step": 1e3 * t / len(secs), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "same_config": whole,
            "config": workload_config(1000), "pairs_per_step_reference": cores if whole else cores * equiv,
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def workload_config(max_iter: int = 1000) -> dict:
    """`config` of BOTH arms (the driver compares them)."""
    return {"workload": "50-frame synthetic 1024x436 sequence (49 pairs), defaults (patch 8, scales 3->0, overlap 0.7, mean-norm, max_iter %d)" % max_iter,
            "mode": "exact (bit-identical to the reference's 1001 passes per patch)", "pairs_per_step_per_gpu": N_PAIRS,
            "l2": "per-step working set ~0.6 GB (pyramids + flows) exceeds the 126 MB L2; no explicit flush"}


def bind_to_gpu_numa(local_rank: int) -> dict:
    """Pin this process to the CPUs of its GPU's NUMA node BEFORE it allocates pinned host memory, so that the staging
    buffers of every rank sit behind their own GPU's root port (without this all ranks of a torchrun job inherit the
    launcher's affinity and their 8 x 175 MB/step of D2H contend on one node)."""
    import torch
    info = {"node": None, "cpus": 0}
    try:
        pr = torch.cuda.get_device_properties(local_rank)
        bus = "%04x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
        node = int(open(f"/sys/bus/pci/devices/{bus}/numa_node").read())
        info["node"] = node
        if node >= 0:
            cpus = set()
            for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
                a, _, b = part.partition("-")
                cpus.update(range(int(a), int(b or a) + 1))
            allowed = os.sched_getaffinity(0)
            use = (cpus & allowed) or cpus
            os.sched_setaffinity(0, use)
            info["cpus"] = len(use)
    except Exception as ex:  # noqa: BLE001
        info["note"] = str(ex)[:120]
    return info


def code_sha():
    """sha256 over the kernel sources: a profile-derived constant (roofline.traffic) is only valid for the code it was
    captured from."""
    import hashlib
    h = hashlib.sha256()
    d = os.path.join(ROOT, "optical-flow-using-dense-inverse-search_b200", "csrc")
    for f in ("dis_kernels.cuh", "dis_search.cu", "dis_pyramid.cu", "dis_densify.cu"):   # the device code of a step (not the host-side files)
        h.update(open(os.path.join(d, f), "rb").read())
    return h.hexdigest()


def parity_check(flows, pairs=(0, 24, 48)):
    """Bit-compare flows of the timed sequence (rank 0: seed 100) with the reference-generated goldens
    (tests/golden/golden_full.npz, made by tests/golden/make_golden_full.py from oracle/_ref).  No oracle code runs."""
    import hashlib
    try:
        z = np.load(os.path.join(ROOT, "tests", "golden", "golden_full.npz"))
        meta = {m["name"]: m for m in json.loads(str(z["meta"]))}
    except Exception as ex:  # noqa: BLE001
        return {"pairs_checked": [], "bit_identical": None, "note": f"goldens unavailable: {ex}"}
    frames, _ = bench_sequence(100)
    ok, checked = True, []
    for i in pairs:
        m = meta.get(f"bench_seq100_pair{i}")
        if m is None:
            continue
        if hashlib.sha256(np.stack([frames[i], frames[i + 1]]).tobytes()).hexdigest() != m["frames_sha"]:
            return {"pairs_checked": checked, "bit_identical": None, "note": "synthetic generator drifted: golden inputs not reproducible here"}
        checked.append(i)
        ok = ok and hashlib.sha256(np.ascontiguousarray(flows[i]).tobytes()).hexdigest() == m["flow_sha"]
    return {"pairs_checked": checked, "bit_identical": bool(ok) if checked else None,
            "against": "tests/golden/golden_full.npz (the reference's own sources, max_iter 1000)"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-fast-mode", action="store_true")
    ap.add_argument("--max-iter", type=int, default=1000)
    ap.add_argument("--no-numa-bind", action="store_true")
    ap.add_argument("--no-config3", action="store_true")
    ap.add_argument("--no-tiled", action="store_true")
    ap.add_argument("--ref-strip", type=int, default=0, help="reference arm: crop pairs to their top H rows (0 = whole pairs)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        reference_arm(args, rank)
        return

    import torch
    import dis_b200
    dist = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    numa = bind_to_gpu_numa(local_rank) if not args.no_numa_bind else {"node": None, "cpus": 0, "note": "--no-numa-bind"}
    frames, gt = bench_sequence(100 + rank)
    eng = dis_b200.Engine(local_rank, max_iter=args.max_iter)
    first = np.arange(N_PAIRS, dtype=np.int32)

    # ---- value: inputs resident in HBM, device time by CUDA events on the engine's stream -------
    eng.upload_frames(frames, first, first + 1)
    for _ in range(args.warmup):
        eng.run_resident()
    eng.synchronize()
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    search_ms, lvl0_ms = 0.0, 0.0
    eng.timer_start()
    for _ in range(args.steps):
        eng.run_resident()
    dev_ms = eng.timer_stop_ms()
    barrier()
    dev_ms = max_over_ranks(dev_ms)
    # per-launch durations of the three search kernels, measured live (events around each launch, last step)
    search_ms, n_search = eng.last_search_ms()
    nl = eng.num_levels()
    lvl_ms = [eng.last_search_level_ms(l) for l in range(nl)]
    phase_ms = [[eng.last_search_phase_ms(l, ph) for l in range(nl)] for ph in range(3)]
    st = eng.stats()
    launches_per_step = st["kernel_launches"]

    # ---- e2e: the public host-buffer call, pinned host memory, H2D + D2H inside the timed region --
    pin_in = torch.empty((N_FRAMES, H, W), dtype=torch.uint8).pin_memory()
    pin_out = torch.empty((N_PAIRS, H, W, 2), dtype=torch.float32).pin_memory()
    pin_in.numpy()[:] = frames
    for _ in range(2):
        eng.run_sequence_ptr(pin_in.data_ptr(), N_FRAMES, W, H, pin_out.data_ptr())
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        eng.run_sequence_ptr(pin_in.data_ptr(), N_FRAMES, W, H, pin_out.data_ptr())
    blk_s = time.perf_counter() - t0
    barrier()
    blk_s = max_over_ranks(blk_s)
    # the same call through the asynchronous form: two engines on this GPU, two submissions in flight, so that the
    # uploads / downloads of one step run under the kernels of the next (every step still copies its inputs up from
    # pinned memory and its flows down; a step's buffers are reused only after its ticket has been waited for)
    pool = dis_b200.Pool([local_rank, local_rank], max_iter=args.max_iter)
    pin_out2 = torch.empty_like(pin_out).pin_memory()
    outs = [pin_out, pin_out2]
    pend = []
    for i in range(4):
        pend.append(pool.submit_sequence_ptr(pin_in.data_ptr(), N_FRAMES, W, H, outs[i % 2].data_ptr()))
        if len(pend) == 2:
            pool.wait(pend.pop(0))
    pool.wait_all(); pend = []
    barrier()
    t0 = time.perf_counter()
    for i in range(args.steps):
        pend.append(pool.submit_sequence_ptr(pin_in.data_ptr(), N_FRAMES, W, H, outs[i % 2].data_ptr()))
        if len(pend) == 2:
            pool.wait(pend.pop(0))
    pool.wait_all()
    e2e_s = time.perf_counter() - t0
    e2e_local_s = e2e_s
    barrier()
    e2e_s = max_over_ranks(e2e_s)
    async_same = bool(np.array_equal(pin_out.numpy().view(np.uint32), pin_out2.numpy().view(np.uint32)))
    if rank == 0 and args.max_iter == 1000:   # what the asynchronous steps left in the host buffers, against the goldens
        async_same = async_same and parity_check(pin_out.numpy())["bit_identical"] and parity_check(pin_out2.numpy())["bit_identical"]
    pool.close()
    del pin_out2
    clocks = sampler.stop() if sampler else None
    checksum = float(np.abs(pin_out.numpy()).mean())
    # host <-> device copy rates of this rank's pinned buffers (informational: where the e2e time goes at N GPUs)
    dev_in = torch.empty_like(pin_in, device="cuda")
    dev_out = torch.empty_like(pin_out, device="cuda")
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    dev_in.copy_(pin_in, non_blocking=True); dev_out.copy_(pin_out, non_blocking=True); torch.cuda.synchronize()
    barrier()
    evs[0].record(); dev_in.copy_(pin_in, non_blocking=True); evs[1].record(); pin_out.copy_(dev_out, non_blocking=True); evs[2].record()
    torch.cuda.synchronize()
    copy_rates = {"rank": rank, "numa": numa, "h2d_gbs": pin_in.numel() / (evs[0].elapsed_time(evs[1]) * 1e6),
                  "d2h_gbs": pin_out.numel() * 4 / (evs[1].elapsed_time(evs[2]) * 1e6), "e2e_ms_per_step": 1e3 * e2e_local_s / args.steps}
    del dev_in, dev_out
    eng.run_sequence_ptr(pin_in.data_ptr(), N_FRAMES, W, H, pin_out.data_ptr())   # restore the flows for the parity check
    if dist is not None:
        all_rates = [None] * world
        dist.all_gather_object(all_rates, copy_rates)
    else:
        all_rates = [copy_rates]

    # parity of what was just timed: the e2e buffer AND the device-resident result against the reference-generated goldens
    parity = None
    if rank == 0 and args.max_iter == 1000:
        parity = parity_check(pin_out.numpy())
        eng.upload_frames(frames, first, first + 1); eng.run_resident()
        res = parity_check(eng.download_flows())
        parity["resident_run_bit_identical"] = res["bit_identical"]

    # ---- BASELINE config 3 as written: a stream of 4096 independent 1024x436 pairs sharded over the N GPUs (STRONG
    #      scaling: total work fixed).  64 distinct synthetic pairs, cycled; every pass recomputes everything. ----------
    config3 = None
    if not args.no_config3:
        C3_TOTAL, C3_CHUNK = 4096, 64
        lo, hi = dis_b200.parallel.shard_pairs(C3_TOTAL // C3_CHUNK, world, rank)     # chunks of 64 pairs, contiguous per rank
        prs = [dis_b200.synth.make_pair(W, H, seed=300 + i)[:2] for i in range(C3_CHUNK)]
        fr = np.concatenate([np.stack([p[0] for p in prs]), np.stack([p[1] for p in prs])])
        idx = np.arange(C3_CHUNK, dtype=np.int32)
        eng.upload_frames(fr, idx, idx + C3_CHUNK)
        eng.run_resident(); eng.synchronize()
        barrier()
        eng.timer_start()
        for _ in range(hi - lo):
            eng.run_resident()
        c3_ms = max_over_ranks(eng.timer_stop_ms())
        barrier()
        config3 = {"workload": "4096 independent synthetic 1024x436 pairs (64 distinct seeds, cycled) sharded over the GPUs in contiguous blocks",
                   "scaling": "strong", "pairs_total": C3_TOTAL, "n_gpus": world, "seconds": c3_ms * 1e-3,
                   "value": C3_TOTAL / (c3_ms * 1e-3), "unit": UNIT, "timing": "device (CUDA events), max over ranks, inputs resident"}

    # ---- informational: two engines on this GPU, both with the timed sequence resident, stepping alternately: the
    #      latency-bound tails of one step's levels are filled by the other's blocks (why e2e can exceed `value`) -------
    two = None
    if rank == 0 and world == 1:
        eb = dis_b200.Engine(local_rank, max_iter=args.max_iter)
        eng.upload_frames(frames, first, first + 1); eb.upload_frames(frames, first, first + 1)
        for e_ in (eng, eb):
            e_.run_resident()
        eng.synchronize(); eb.synchronize()
        k2 = max(2, args.steps // 2 * 2)
        t0 = time.perf_counter()
        for i in range(k2):
            (eng if i % 2 == 0 else eb).run_resident()       # asynchronous: both streams stay fed
        eng.synchronize(); eb.synchronize()
        dt = time.perf_counter() - t0
        two = {"value": N_PAIRS * k2 / dt, "unit": UNIT, "ms_per_step": 1e3 * dt / k2, "steps": k2,
               "timing": "host clock around device-resident steps issued alternately to two engines (no copies)"}
        eb.close()

    # ---- informational: the optional convergence early-out (conv_eps > 0), a REPORTED DEVIATION from the
    #      reference (which has no convergence test); never the headline ------------------------
    fast = None
    if rank == 0 and world == 1 and not args.no_fast_mode:
        exact = pin_out.numpy()           # the exact flows of this sequence (e2e buffer, checked against the goldens above)
        fe = dis_b200.Engine(local_rank, max_iter=args.max_iter, conv_eps=1e-4)
        fe.upload_frames(frames, first, first + 1)
        for _ in range(3):
            fe.run_resident()
        fe.synchronize()
        fe.timer_start()
        for _ in range(args.steps):
            fe.run_resident()
        fms = fe.timer_stop_ms()
        ff = fe.download_flows()
        fst = fe.stats()
        fe.close()
        fast = {"conv_eps": 1e-4, "value": N_PAIRS * args.steps / (fms * 1e-3), "unit": UNIT,
                "mean_end_point_difference_vs_exact_px": float(np.linalg.norm(ff - exact, axis=-1).mean()),
                "epe_vs_ground_truth_px": {"exact": float(np.linalg.norm(exact - gt, axis=-1).mean()),
                                           "fast": float(np.linalg.norm(ff - gt, axis=-1).mean())},
                "patch_iterations_per_step": fst["patch_iterations"],
                "note": "stops a patch once |delta_u| < conv_eps; within the 0.01 px dense gate, misses the 99.9 % per-patch gate (SURVEY.md App. B)"}

    # ---- BASELINE config 5: ONE 3840x2160 pair split over the N GPUs by bands of grid rows, halo exchange over NVLink
    #      peer memory (dis_pool_run_pair_tiled_u8).  In-process: rank 0 drives all N GPUs, the other ranks wait. -----------
    tiled = None
    if world > 1 and not args.no_tiled:
        # The other ranks must not wait INSIDE an NCCL collective while rank 0 drives their GPUs: a pending NCCL barrier is
        # a kernel spinning on each of those GPUs (measured: it turns the 1.5x of two GPUs into 0.8x).  They wait on the CPU.
        cpu_group = dist.new_group(backend="gloo")
        barrier()
        if rank == 0:
            try:
                a4, b4, _ = dis_b200.synth.make_pair(3840, 2160, seed=500, amp=11.0)
                one = dis_b200.Engine(local_rank, max_iter=args.max_iter)
                pa = torch.from_numpy(a4).pin_memory(); pb = torch.from_numpy(b4).pin_memory()
                po1 = torch.empty((2160, 3840, 2), dtype=torch.float32).pin_memory()
                poN = torch.empty((2160, 3840, 2), dtype=torch.float32).pin_memory()
                lib = dis_b200.load_library()
                import ctypes as C

                def single():
                    rc = lib.dis_run_pair_u8(one._h, C.c_void_p(pa.data_ptr()), C.c_void_p(pb.data_ptr()), 3840, 2160, 3840, C.c_void_p(po1.data_ptr()))
                    assert rc == 0
                single()
                t0 = time.perf_counter()
                for _ in range(3):
                    single()
                one_ms = (time.perf_counter() - t0) / 3 * 1e3
                one.upload_frames(np.stack([a4, b4]), np.array([0], np.int32), np.array([1], np.int32))
                one.run_resident(); one.synchronize(); one.timer_start()
                for _ in range(3):
                    one.run_resident()
                one_dev_ms = one.timer_stop_ms() / 3
                one.close()
                pool = dis_b200.Pool(list(range(world)), max_iter=args.max_iter)
                pool.run_pair_tiled_ptr(pa.data_ptr(), pb.data_ptr(), 3840, 2160, poN.data_ptr())
                t0 = time.perf_counter()
                dev = []
                for _ in range(3):
                    pool.run_pair_tiled_ptr(pa.data_ptr(), pb.data_ptr(), 3840, 2160, poN.data_ptr())
                    dev.append(pool.last_tiled_device_ms())
                n_ms = (time.perf_counter() - t0) / 3 * 1e3
                pool.close()
                tiled = {"workload": "one synthetic 3840x2160 pair, defaults, max_iter %d, bands of grid rows + per-level halo exchange over peer memory" % args.max_iter,
                         "n_gpus": world, "ms_per_pair_host_to_host": n_ms, "ms_per_pair_device": float(np.mean(dev)),
                         "one_gpu_ms_per_pair_host_to_host": one_ms, "one_gpu_ms_per_pair_device": one_dev_ms,
                         "speedup_host_to_host": one_ms / n_ms, "speedup_device": one_dev_ms / float(np.mean(dev)),
                         "bit_identical_to_one_gpu": bool(np.array_equal(po1.numpy().view(np.uint32), poN.numpy().view(np.uint32)))}
            except Exception as ex:  # noqa: BLE001
                tiled = {"error": str(ex)[:300]}
            torch.cuda.set_device(local_rank)
        dist.barrier(group=cpu_group)
        barrier()

    # run counters summed over ranks (NCCL all-reduce; statistics only — no pixel data crosses ranks)
    job = dis_b200.parallel.reduce_stats(st, dev_ms, dist, "cuda")
    if rank == 0:
        hbm_peak, sm_max, peak_src = measured_peaks()
        value = world * N_PAIRS * args.steps / (dev_ms * 1e-3)
        e2e = world * N_PAIRS * args.steps / e2e_s
        infos = [eng.level_info(l) for l in range(nl)]
        patches = sum(li["num_patches"] for li in infos) * N_PAIRS
        conv_ms, walk_ms, strag_ms = (sum(x) for x in phase_ms)
        step_ms = dev_ms / args.steps
        sm_mhz = (clocks or {}).get("sm_mhz") or sm_max
        n_sm = torch.cuda.get_device_properties(local_rank).multi_processor_count
        # ---- the roof of this path is fp32 instruction issue WITHOUT FMA, not HBM: executed work =
        #      sampled passes x (13 ps^2 + 32 = 864 lane-ops) + replayed (memoised) passes x 11 lane-ops, over the time
        #      the three search kernels of the four levels take; peak = SMs x 128 fp32 lanes x the clock measured live.
        issue_peak = n_sm * 128 * sm_mhz * 1e6
        sampled = st["sampled_iterations"]
        replayed = st["patch_iterations"] - sampled
        s_conv, s_walk = st["sampled_converge"], st["sampled_walk"]
        s_strag = sampled - s_conv - s_walk
        lane_ops = sampled * F_ITER + replayed * 11
        achieved = lane_ops / (search_ms * 1e-3)
        # ---- HBM: 26.0 MB compulsory per Sintel pair (inputs once, four live planes written + read
        #      once, level flows written once and the coarser ones read once)
        HBM_PER_PAIR = 26.0e6
        hbm_achieved = HBM_PER_PAIR * N_PAIRS / (step_ms * 1e-3) / 1e9
        traffic, traffic_note = None, None
        tpath = os.path.join(ROOT, "profiles", "r2_search_traffic.json")   # dram bytes of the search launches of one step: ncu --set full
        if os.path.exists(tpath):
            tj = json.load(open(tpath))
            if tj.get("code_sha") == code_sha():
                traffic = tj.get("dram_bytes_per_step")
            else:
                traffic_note = "profiles/r2_search_traffic.json was captured from other kernel sources (code_sha differs): not reported"
        else:
            traffic_note = "no ncu capture on file for these kernel sources"
        it_rate = st["patch_iterations"] / (search_ms * 1e-3)
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": step_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": workload_config(args.max_iter),
            "mpix_per_s": value * W * H / 1e6,
            "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": int(pin_in.numel()), "d2h_bytes_per_step": int(pin_out.numel() * 4),
                    "api": "dis_pool_submit_sequence_u8 + dis_pool_wait on a pool of two engines on this GPU, two steps in flight (pinned host buffers)",
                    "ms_per_step": 1e3 * e2e_s / args.steps, "bit_identical_to_goldens_and_between_engines": async_same,
                    "blocking_call": {"api": "dis_run_sequence_u8, one engine, one call at a time", "value": world * N_PAIRS * args.steps / blk_s,
                                      "ms_per_step": 1e3 * blk_s / args.steps}},
            "gpu_launches": int(launches_per_step * args.steps),
            "parity": parity,
            "roofline": {"kernel": "the inverse search = k_converge + k_inverse_search (memoised walk) + k_stragglers, 4 levels (figures per step)",
                         "bound": "fp32_issue", "achieved": achieved / 1e12, "peak": issue_peak / 1e12, "unit": "Tlane-op/s (fp32, no FMA)",
                         "frac": achieved / issue_peak, "traffic": traffic, "traffic_note": traffic_note,
                         "peak_source": f"{n_sm} SMs x 128 fp32 lanes x {sm_mhz:.0f} MHz (median SM clock sampled during the timed region)",
                         "executed": {"sampled_passes": sampled, "replayed_passes": replayed, "lane_ops_per_sampled_pass": F_ITER,
                                      "lane_ops_per_replayed_pass": 11},
                         "kernel_ms_per_step": search_ms,
                         "reference_equivalent_frac": it_rate * F_ITER / issue_peak,
                         "note": "reference_equivalent_frac counts all 1001 passes per patch as if sampled: work skipped by exact memoisation, "
                                 "NOT a fraction of peak"},
            "roofline_hbm": {"bound": "hbm", "achieved": hbm_achieved, "peak": hbm_peak, "unit": "GB/s", "frac": hbm_achieved / hbm_peak,
                             "algorithmic_bytes_per_pair": HBM_PER_PAIR, "peak_source": peak_src + " (burst figure of MEASURED_PEAKS.json)",
                             "note": "SURVEY.md 8(d)-i compulsory bytes over the whole step; the path is not HBM-bound"},
            "roofline_kernels": [
                {"kernel": "k_converge<8,4>", "ms_per_step": conv_ms, "sampled_passes": s_conv,
                 "fp32_lane_ops_per_s": s_conv * F_ITER / (conv_ms * 1e-3), "frac_fp32_issue": s_conv * F_ITER / (conv_ms * 1e-3) / issue_peak,
                 "ms_per_level_fine_to_coarse": phase_ms[0]},
                {"kernel": "k_inverse_search<8,8> (walk)", "ms_per_step": walk_ms, "sampled_passes": s_walk, "replayed_passes": replayed,
                 "fp32_lane_ops_per_s": (s_walk * F_ITER + replayed * 11) / (walk_ms * 1e-3),
                 "frac_fp32_issue": (s_walk * F_ITER + replayed * 11) / (walk_ms * 1e-3) / issue_peak,
                 "ms_per_level_fine_to_coarse": phase_ms[1]},
                {"kernel": "k_stragglers<8,4>", "ms_per_step_not_hidden_by_the_walk": strag_ms, "sampled_passes": s_strag,
                 "ms_per_level_fine_to_coarse": phase_ms[2]}],
            "config3_4096_pairs": config3, "tiled_4k": tiled, "two_engines_resident_informational": two, "per_rank_copy_rates": all_rates,
            "stats_per_step": st, "stats_job_per_step": job, "search_share_of_step": search_ms / step_ms,
            "clocks": clocks, "checksum": checksum,
        }
        if fast is not None:
            line["fast_mode_informational"] = fast
        if world == 1 and not args.no_cpu_baseline:
            import oracle
            kind = "reference" if oracle.load_ref() is not None else "port"
            v, dt, equiv = cpu_run(kind, 1, 216, first_pair=0)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": 1, "kind": kind,
                                    "sample": f"pair 0 of the timed sequence cropped to 1024x216 (= {equiv:.4f} of a 1024x436 pair by patch count), "
                                              f"single thread, {dt:.1f} s; --impl reference times whole pairs on all cores"}
        print(json.dumps(line), flush=True)
    eng.close()
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
