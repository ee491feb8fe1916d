#!/usr/bin/env python
"""Writes one compact SASS listing per kernel of liborbx_b200.so into profiles/sass/ (encodings
stripped).  usage: python profiles/dump_sass.py [round_tag]"""
import os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1] if len(sys.argv) > 1 else "r1"
KERNELS = ["k_resize_rows_async", "k_resize_rows", "k_resize_level", "k_fast_cells", "k_octree",
           "k_describe", "k_hamming_csr", "k_knn2", "k_stereo_rows", "k_stereo_refine", "k_grid_build", "k_area",
           "k_area_onepass", "k_rowptr_scan", "k_input", "k_resize_tail", "k_undistort", "k_scatter", "k_distinctive", "k_bow_descend", "k_bow_topk", "k_csr_topk"]
txt = subprocess.run(["cuobjdump", "-sass", os.path.join(ROOT, "orb-slam3-mac_b200", "liborbx_b200.so")],
                     capture_output=True, text=True).stdout
os.makedirs(os.path.join(ROOT, "profiles", "sass"), exist_ok=True)
for part in re.split(r"(?=\t\tFunction : )", txt)[1:]:
    mangled = re.match(r"\t\tFunction : (\S+)", part).group(1)
    name = next((k for k in KERNELS if re.search(r"\d+" + k + r"(?![a-z_])", mangled)), None)
    if not name:
        continue
    lines = [re.sub(r"\s*/\* 0x[0-9a-f]+ \*/\s*$", "", l) for l in part.splitlines()
             if not re.match(r"^\s*/\* 0x[0-9a-f]+ \*/\s*$", l)]
    with open(os.path.join(ROOT, "profiles", "sass", f"{tag}_{name}.sass"), "w") as f:
        f.write("\n".join(lines) + "\n")
    ops = {}
    for l in lines:
        m = re.match(r"\s*/\*[0-9a-f]{4}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", l)
        if m:
            ops[m.group(1)] = ops.get(m.group(1), 0) + 1
    top = ", ".join(f"{k} {v}" for k, v in sorted(ops.items(), key=lambda x: -x[1])[:8])
    print(f"{name:22s} {sum(ops.values()):5d} instr: {top}")
