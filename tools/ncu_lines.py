"""Per-CUDA-source-line instruction and stall attribution from an ncu report.

usage: python tools/ncu_lines.py <report.ncu-rep> [kernel substring] [top N]
Reads `ncu --page source --print-source cuda,sass --csv` (needs -lineinfo and --import-source on)
and, for the launch of the matching kernel with the most executed instructions, prints the hottest
source lines: share of warp instructions, share of stall samples, top stall reasons.
"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
kname = sys.argv[2] if len(sys.argv) > 2 else ""
topn = int(sys.argv[3]) if len(sys.argv) > 3 else 40
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
secs = [i for i, r in enumerate(rows) if r and r[0] == "Function Name"]
best = None
for k, i in enumerate(secs):
    if kname not in rows[i][1]:
        continue
    end = secs[k + 1] - 1 if k + 1 < len(secs) else len(rows)
    hdr, body = rows[i + 1], rows[i + 2:end]
    ci, cs = hdr.index("Instructions Executed"), hdr.index("# Samples")
    cu = [r for r in body if len(r) > ci and r[2] == "-" and r[0].isdigit()]
    tot = sum(float(r[ci] or 0) for r in cu)
    if best is None or tot > best[0]:
        best = (tot, rows[i - 1][1], rows[i][1], hdr, cu)
if best is None:
    sys.exit("kernel not found")
tot, path, fn, hdr, cu = best
ci, cs = hdr.index("Instructions Executed"), hdr.index("# Samples")
stall = [j for j, h in enumerate(hdr) if h.startswith("stall_")]
ts = sum(float(r[cs] or 0) for r in cu)
print(f"{fn}\n{path}: {tot:.4g} warp instructions, {ts:.0f} samples")
for r in sorted(cu, key=lambda r: -float(r[ci] or 0))[:topn]:
    a = float(r[cs] or 0)
    top = sorted(((float(r[j] or 0), hdr[j][6:]) for j in stall), reverse=True)[:3]
    st = " ".join(f"{n}={100 * v / a:.0f}%" for v, n in top if a and v)
    print(f"{int(r[0]):4d} inst {100 * float(r[ci] or 0) / tot:5.1f}% smp {100 * a / ts:5.1f}% | {r[1].strip()[:70]:70s} | {st}")
