"""Development loop helper (GPU box): run the bench workload (49-pair sequence, device-resident) once per
environment given on the command line, each in its own process, and print pairs/s, phase times and a sha256 of
all flows (so that a tuning variant that changes results is caught at once).
usage: python tools/exp.py [--reps N] [--frames F] "NAME: ENV1=a ENV2=b" "NAME2: DIS_B200_LIB=tools/variants/x.so" ...
       python tools/exp.py --child   (internal)"""
import hashlib
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def child(reps, n_frames):
    import numpy as np
    import dis_b200
    if os.environ.get("EXP_HQ"):      # BASELINE config 4: 1920x1080, patch 12, overlap 0.75, scales 5->0; n_frames - 1 independent pairs
        n = n_frames - 1
        prs = [dis_b200.synth.make_pair(1920, 1080, seed=400 + i, amp=5.6)[:2] for i in range(n)]
        frames = np.concatenate([np.stack([p[0] for p in prs]), np.stack([p[1] for p in prs])])
        e = dis_b200.Engine(0, patch_size=12, patch_overlap=0.75, coarsest_scale=5)
        first = np.arange(n, dtype=np.int32)
        e.upload_frames(frames, first, first + n)
    else:
        frames, _ = dis_b200.synth.make_sequence(n_frames, 1024, 436, seed=100)
        e = dis_b200.Engine(0)
        first = np.arange(n_frames - 1, dtype=np.int32)
        e.upload_frames(frames, first, first + 1)
    e.run_resident(); e.synchronize()
    e.timer_start()
    for _ in range(reps):
        e.run_resident()
    ms = e.timer_stop_ms() / reps
    st = e.stats()
    nl = e.num_levels()
    ph = [[round(e.last_search_phase_ms(l, p), 2) for l in range(nl)] for p in range(3)]
    sha = hashlib.sha256(e.download_flows().tobytes()).hexdigest()[:16]
    print(f"{(n_frames - 1) / ms * 1e3:8.1f} pairs/s  step {ms:6.2f} ms | conv {ph[0]} {sum(ph[0]):.2f} | walk {ph[1]} {sum(ph[1]):.2f} | strag {ph[2]} {sum(ph[2]):.2f} | "
          f"samples/patch {st['sampled_iterations'] / st['patches']:.2f} | sha {sha}", flush=True)
    e.close()


if __name__ == "__main__":
    args = sys.argv[1:]
    reps, frames = 3, 50
    if "--reps" in args:
        i = args.index("--reps"); reps = int(args[i + 1]); del args[i:i + 2]
    if "--frames" in args:
        i = args.index("--frames"); frames = int(args[i + 1]); del args[i:i + 2]
    if args and args[0] == "--child":
        child(reps, frames)
        sys.exit(0)
    for spec in args or ["base:"]:
        name, _, envs = spec.partition(":")
        env = dict(os.environ)
        for kv in envs.split():
            k, _, v = kv.partition("=")
            env[k] = v
        try:
            out = subprocess.run([sys.executable, __file__, "--reps", str(reps), "--frames", str(frames), "--child"], env=env, capture_output=True, text=True, timeout=600)
            line = out.stdout.strip().splitlines()[-1] if out.stdout.strip() else ("FAILED rc=%d %s" % (out.returncode, out.stderr.strip()[-300:]))
        except subprocess.TimeoutExpired:
            line = "TIMEOUT"
        print(f"{name.strip():>28s}: {line}", flush=True)
