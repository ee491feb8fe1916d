#!/usr/bin/env python
"""Condensed per-launch table from an ncu --set full report.  usage: ncu_summary.py report.ncu-rep"""
import csv, io, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units = rows[0], rows[1]
cols = [("Kernel Name", "kernel"), ("gpu__time_duration.sum", "time"), ("dram__bytes_read.sum", "dram_rd"),
        ("dram__bytes_write.sum", "dram_wr"), ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
        ("launch__registers_per_thread", "regs"), ("smsp__inst_executed.sum", "warp_inst"),
        ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem_conflicts"),
        ("l1tex__t_sector_hit_rate.pct", "l1hit%"), ("lts__t_sector_hit_rate.pct", "l2hit%"),
        ("smsp__thread_inst_executed_per_inst_executed.ratio", "thr/inst")]
idx = [(hdr.index(c), n) for c, n in cols if c in hdr]
print("| " + " | ".join(n + (f" [{units[i]}]" if units[i] else "") for i, n in idx) + " |")
print("|" + "---|" * len(idx))
for r in rows[2:]:
    vals = []
    for i, n in idx:
        v = r[i]
        if n == "kernel":
            v = v.split("(")[0]
        else:
            try:
                v = f"{float(v.replace(',', '')):.4g}"
            except ValueError:
                pass
        vals.append(v)
    print("| " + " | ".join(vals) + " |")
