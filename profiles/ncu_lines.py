#!/usr/bin/env python
"""Per-source-line instruction / stall-sample totals from an ncu report (needs -lineinfo).
usage: ncu_lines.py report.ncu-rep kernel_regex [top_n]"""
import csv, io, subprocess, sys

rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass",
                      "--kernel-name", "regex:" + kern], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
data, hdr, done = [], None, False
for r in rows:
    if len(r) > 8 and r[0] == "Line No":
        if hdr is not None:
            break  # only the first launch of the kernel
        hdr = r
        ii, ss, ta = r.index("Instructions Executed"), r.index("# Samples"), r.index("Avg. Threads Executed")
        continue
    if hdr is None or len(r) <= ii or not r[0].isdigit():
        continue
    try:
        data.append((int(r[ii]), int(r[ss]), int(r[0]), r[1].strip(), r[ta]))
    except ValueError:
        pass
tot = sum(d[0] for d in data) or 1
tots = sum(d[1] for d in data) or 1
print(f"{kern}: total warp-instructions {tot}, stall samples {tots}")
for n, s, ln, src, ta in sorted(data, reverse=True)[:top]:
    print(f"{n:11d} {100*n/tot:5.1f}%  smp {100*s/tots:5.1f}%  thr {ta:>4}  L{ln:>4} | {src[:96]}")
