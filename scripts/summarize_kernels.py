#!/usr/bin/env python
"""gpurun_out/kernels_<tag>.csv (scripts/r02_ncu_kernels.sh) -> profiles/<tag>_kernels.md: per kernel the average device
time, DRAM read + write bytes per launch and, where DESIGN.md states algorithmic bytes, achieved GB/s and fractions.
  python scripts/summarize_kernels.py r02_1m 1000000 318318        (tag, bodies, nodes)"""
import collections
import csv
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag, n, nn = sys.argv[1], float(sys.argv[2]), float(sys.argv[3])
peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
lines = [l for l in open(os.path.join(ROOT, "gpurun_out", f"kernels_{tag}.csv")) if not l.startswith("==")]
agg = collections.OrderedDict()
for row in csv.DictReader(lines):
    k = row["Kernel Name"].split("(")[0].replace("void ", "").replace("nb::", "")
    if k.startswith("fp32_ubench"):
        continue
    v = float(row["Metric Value"].replace(",", ""))
    u = row["Metric Unit"]
    v *= {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1, "ns": 1, "us": 1e3, "ms": 1e6}.get(u, 1)
    agg.setdefault(k, collections.defaultdict(list))[row["Metric Name"]].append(v)
alg = {"morton_repair_kernel": 29 * n, "sort_repair_kernel": 21 * n, "morton_kernel": 20 * n, "morton_hist_kernel": 20 * n, "sort_histogram_kernel": 8 * n, "sort_onesweep_kernel": 24 * n, "tree_count_kernel": 13 * n,
       "tree_emit_quad_kernel": 84 * nn, "tree_link_kernel": 56 * nn, "tree_finish_kernel": 40 * n + 100 * nn,
       "integrate_resort_kernel": 60 * n}
with open(os.path.join(ROOT, "profiles", f"{tag}_kernels.md"), "w") as f:
    f.write(f"# {tag}: every kernel of the step under ncu (gpu__time_duration, dram__bytes_read + write; --clock-control none)\n\n"
            f"`bash scripts/r02_ncu_kernels.sh {tag} ...`; {int(n)} bodies, {int(nn)} nodes.  Times under ncu are cold-cache and\n"
            "serialised (compare shares; the CUDA-event times of the same kernels are in the bench line's `roofline_hbm`).\n"
            f"Algorithmic bytes as in DESIGN.md §4; peak = {peak:.0f} GB/s (MEASURED_PEAKS.json).\n\n"
            "| kernel | launches | avg us | DRAM MB / launch | algorithmic MB | DRAM / algorithmic | algorithmic GB/s | frac of peak |\n"
            "|---|---:|---:|---:|---:|---:|---:|---:|\n")
    for k, m in sorted(agg.items(), key=lambda kv: -sum(kv[1]["gpu__time_duration.sum"])):
        t = m["gpu__time_duration.sum"]
        d = [a + b for a, b in zip(m["dram__bytes_read.sum"], m["dram__bytes_write.sum"])]
        us, mb = sum(t) / len(t) / 1e3, sum(d) / len(d) / 1e6
        base = k.split("<")[0]
        if base in alg:
            a = alg[base]
            f.write(f"| `{k}` | {len(t)} | {us:.1f} | {mb:.1f} | {a / 1e6:.1f} | {mb * 1e6 / a:.2f} | {a / (us * 1e-6) / 1e9:.0f} | {a / (us * 1e-6) / 1e9 / peak:.3f} |\n")
        else:
            f.write(f"| `{k}` | {len(t)} | {us:.1f} | {mb:.1f} | | | | |\n")
print("wrote", f"profiles/{tag}_kernels.md")
