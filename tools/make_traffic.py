"""profiles/r2_search_traffic.json from an ncu launch list (csv with gpu__time_duration.sum, dram__bytes_read.sum,
dram__bytes_write.sum per launch, taken with tools/prof_run.py on the 49-pair bench workload): DRAM bytes and times of
the kernels of ONE step, stamped with the sha256 of the kernel sources so that bench.py can refuse a stale file.
usage: python tools/make_traffic.py gpurun_out/launches.csv [out.json]"""
import csv
import json
import os
import subprocess
import sys
from collections import defaultdict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

rows = list(csv.DictReader(l for l in open(sys.argv[1]) if l.startswith('"')))
launch = defaultdict(dict)
for r in rows:
    launch[int(r["ID"])]["name"] = r["Kernel Name"].split("(")[0].replace("void ", "").replace("unnamed>::", "")
    launch[int(r["ID"])][r["Metric Name"]] = float(r["Metric Value"].replace(",", ""))
ids = sorted(launch)
# one step = from a level-0 pyramid launch (k_level<0/1>) to the launch before the next one
starts = [i for i in ids if launch[i]["name"].startswith(("k_level<0", "k_level<1"))]
assert starts, "the capture must contain the start of a step (the level-0 pyramid launch)"
end = starts[1] if len(starts) > 1 else ids[-1] + 1
step = [i for i in ids if starts[0] <= i < end]
assert any(launch[i]["name"].startswith("k_crop") or launch[i]["name"].startswith("k_upsample") for i in step), "the capture must contain one whole step"
per = defaultdict(lambda: {"launches": 0, "ms": 0.0, "dram_read": 0.0, "dram_write": 0.0})
for i in step:
    L = launch[i]
    k = L["name"].split("<")[0]
    per[k]["launches"] += 1
    per[k]["ms"] += L.get("gpu__time_duration.sum", 0.0) / 1e6
    per[k]["dram_read"] += L.get("dram__bytes_read.sum", 0.0)
    per[k]["dram_write"] += L.get("dram__bytes_write.sum", 0.0)
search = [k for k in per if k in ("k_converge", "k_inverse_search", "k_stragglers")]
out = {"source": os.path.basename(sys.argv[1]) + " (ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none; launches serialised, cold cache)",
       "workload": "tools/prof_run.py: the 49-pair bench sequence, device-resident, one step",
       "code_sha": bench.code_sha(),
       "git_head": subprocess.run(["git", "-C", ROOT, "rev-parse", "HEAD"], capture_output=True, text=True).stdout.strip(),
       "dram_bytes_per_step": sum(per[k]["dram_read"] + per[k]["dram_write"] for k in search),
       "dram_bytes_per_step_all_kernels": sum(v["dram_read"] + v["dram_write"] for v in per.values()),
       "kernels": {k: v for k, v in per.items()}}
dst = sys.argv[2] if len(sys.argv) > 2 else os.path.join(ROOT, "profiles", "r2_search_traffic.json")
json.dump(out, open(dst, "w"), indent=1)
print(json.dumps(out, indent=1))
