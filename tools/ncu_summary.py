"""Compact per-launch table from an ncu report (development aid; the outputs under profiles/ come from it).
usage: python tools/ncu_summary.py report.ncu-rep"""
import csv
import subprocess
import sys

txt = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
hdr, units = rows[0], rows[1]
cols = [("Kernel Name", "kernel"), ("Grid Size", "grid"), ("Block Size", "block"), ("launch__registers_per_thread", "regs"),
        ("gpu__time_duration.sum", "time"), ("smsp__inst_executed.sum", "warp_inst"),
        ("sm__inst_executed.avg.per_cycle_elapsed", "ipc"), ("smsp__thread_inst_executed_per_inst_executed.ratio", "lanes/inst"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occupancy%"),
        ("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "fma_pipe%"),
        ("l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "l1_data_pipe%"),
        ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "smem_wavefronts%"),
        ("lts__t_sector_hit_rate.pct", "l2_hit%"), ("dram__bytes_read.sum", "dram_read"), ("dram__bytes_write.sum", "dram_write"),
        ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
        ("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "stall_long_sb"),
        ("smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "stall_short_sb"),
        ("smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "stall_wait"),
        ("smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio", "stall_not_selected"),
        ("smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio", "stall_branch")]
idx = [(hdr.index(k), n) for k, n in cols if k in hdr]
print("\t".join(f"{n}[{units[i]}]" if units[i] else n for i, n in idx))
for r in rows[2:]:
    out = []
    for i, n in idx:
        v = r[i]
        if n == "kernel":
            v = v.split("(")[0].replace("void ", "").replace("unnamed>::", "")
        else:
            try:
                v = f"{float(v):.4g}"
            except ValueError:
                pass
        out.append(v)
    print("\t".join(out))
