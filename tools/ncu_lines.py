"""Attribute ncu per-SASS-instruction counters to CUDA source lines.

usage: python tools/ncu_lines.py <report.ncu-rep> <libdis_b200.so> <kernel substring> [launch index]
Joins `ncu --page source --csv` (SASS view: executed instructions, stall samples) with the line table
of the cubin (`nvdisasm -g`) by instruction order, and prints the hottest source lines.
"""
import csv, io, os, re, subprocess, sys, tempfile

rep, so, kname = sys.argv[1], sys.argv[2], sys.argv[3]
launch = int(sys.argv[4]) if len(sys.argv) > 4 else -1
tmp = tempfile.mkdtemp()
subprocess.check_call(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, stdout=subprocess.DEVNULL)
lines = None
for f in os.listdir(tmp):
    out = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, f)], capture_output=True, text=True).stdout
    if kname not in out:
        continue
    # split per function
    cur, fn, table = None, None, {}
    for ln in out.splitlines():
        m = re.match(r"\s*\.text\.(\S+):", ln)
        if m: fn = m.group(1); table[fn] = []; cur = None; continue
        m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
        if m: cur = (os.path.basename(m.group(1)), int(m.group(2))); continue
        if fn and re.match(r"\s+/\*[0-9a-f]{4,}\*/", ln): table[fn].append(cur)
    for k, v in table.items():
        if kname in k and v: lines = v
if lines is None: sys.exit("kernel not found in cubins")
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
secs = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"]
secs = [i for i in secs if kname in rows[i][1] or True]
i = secs[launch]
end = secs[secs.index(i) + 1] if secs.index(i) + 1 < len(secs) else len(rows)
hdr = rows[i + 1]; body = rows[i + 2:end]
ci, cs = hdr.index("Instructions Executed"), hdr.index("# Samples")
stall_cols = [j for j, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
agg = {}
n = min(len(body), len(lines))
if len(body) != len(lines): print(f"warning: {len(body)} SASS rows vs {len(lines)} in cubin", file=sys.stderr)
tot = tots = 0
for r, l in zip(body[:n], lines[:n]):
    a = agg.setdefault(l, [0.0, 0.0, {}])
    a[0] += float(r[ci] or 0); a[1] += float(r[cs] or 0)
    for j in stall_cols:
        v = float(r[j] or 0)
        if v: a[2][hdr[j]] = a[2].get(hdr[j], 0) + v
    tot += float(r[ci] or 0); tots += float(r[cs] or 0)
print(f"kernel launch {launch}: {tot:.4g} warp instructions, {tots:.0f} samples, {n} SASS instructions")
src = {}
for l, a in sorted(agg.items(), key=lambda kv: -kv[1][1])[:45]:
    top = sorted(a[2].items(), key=lambda kv: -kv[1])[:3]
    print(f"{str(l):28s} inst {100*a[0]/tot:5.1f}%  samples {100*a[1]/tots:5.1f}%  " + " ".join(f"{k[6:]}={100*v/a[1]:.0f}%" for k, v in top))
