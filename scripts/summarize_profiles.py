#!/usr/bin/env python
"""Turn gpurun_out/launches[_<tag>].csv and gpurun_out/prof_[<tag>_]<kernel>.ncu-rep into the committed summaries
under profiles/ (run here, no GPU needed):  python scripts/summarize_profiles.py r01c traverse_fast_kernel [more kernels]
The first kernel named is the one bench.py's roofline.traffic is taken from."""
import collections
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1]
kernels = sys.argv[2:] or ["traverse_kernel"]
out_dir = os.path.join(ROOT, "profiles")
os.makedirs(out_dir, exist_ok=True)

# ---- launch list -------------------------------------------------------------------------------
path = os.path.join(ROOT, "gpurun_out", f"launches_{tag}.csv")
if not os.path.exists(path):
    path = os.path.join(ROOT, "gpurun_out", "launches.csv")
lines = [l for l in open(path) if not l.startswith("==")]
agg = collections.OrderedDict()
for row in csv.DictReader(lines):
    k = row["Kernel Name"].split("(")[0].replace("void ", "").replace("nb::", "")
    if k.startswith("fp32_ubench_kernel"):
        continue  # roofline-denominator microbenchmarks bench.py runs after the timed region: not part of a step
    agg.setdefault(k, []).append(float(row["Metric Value"].replace(",", "")))
tot = sum(sum(v) for v in agg.values())
with open(os.path.join(out_dir, f"{tag}_launches.md"), "w") as f:
    f.write(f"# {tag}: ncu launch list (gpu__time_duration.sum, --clock-control none)\n\n"
            "Command: `python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e` (1M-body uniform disc, quadtree,\n"
            "theta 0.5; includes the counter replay and warm-up steps).  Times are cold-cache and serialised: compare SHARES.\n\n"
            "| kernel | launches | avg us | total us | share |\n|---|---:|---:|---:|---:|\n")
    for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
        f.write(f"| `{k}` | {len(v)} | {sum(v) / len(v) / 1e3:.1f} | {sum(v) / 1e3:.1f} | {sum(v) / tot * 100:.1f}% |\n")
print("wrote launches")

# ---- full capture of the top kernel ---------------------------------------------------------------
for kernel in kernels:
    rep = os.path.join(ROOT, "gpurun_out", f"prof_{tag}_{kernel}.ncu-rep")
    if not os.path.exists(rep):
        rep = os.path.join(ROOT, "gpurun_out", f"prof_{kernel}.ncu-rep")
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    want = ["Kernel Name", "gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
            "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
            "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
            "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.sum",
            "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__warps_eligible.avg.per_cycle_active",
            "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__cycles_elapsed.avg",
            "smsp__sass_thread_inst_executed_op_ffma_pred_on.sum", "smsp__sass_thread_inst_executed_op_fadd_pred_on.sum",
            "smsp__sass_thread_inst_executed_op_fmul_pred_on.sum"]
    summary = {}
    with open(os.path.join(out_dir, f"{tag}_{kernel}_ncu.md"), "w") as f:
        f.write(f"# {tag}: ncu --set full, `{kernel}` (one launch inside bench.py, 1M-body disc, quadtree, theta 0.5)\n\n"
                "| metric | unit | value |\n|---|---|---|\n")
        for w in want:
            if w in hdr:
                i = hdr.index(w)
                val = data[0][i]
                summary[w] = (val, units[i])
                f.write(f"| `{w}` | {units[i]} | {val[:90]} |\n")

        def num(name):
            v, u = summary[name]
            x = float(v.replace(",", ""))
            return x * {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1}.get(u, 1)

        if "dram__bytes_read.sum" in summary:
            traffic = num("dram__bytes_read.sum") + num("dram__bytes_write.sum")
            f.write(f"\nDRAM traffic per launch: {traffic / 1e6:.1f} MB (read + write).\n")
            if kernel == kernels[0]:
                json.dump({"dram_bytes_per_launch": traffic, "source": f"profiles/{tag}_{kernel}_ncu.md"},
                          open(os.path.join(out_dir, "traverse_traffic.json"), "w"))
        # per-SASS-instruction execution counts, aggregated
        src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
        srows = list(csv.reader(src.splitlines()))
        sh = srows[1]
        isrc, iex, ith = sh.index("Source"), sh.index("Instructions Executed"), sh.index("Avg. Predicated-On Threads Executed")
        ops = collections.Counter()
        total = 0
        lanes = 0.0
        for r in srows[2:]:
            n = int(r[iex])
            total += n
            lanes += n * float(r[ith])
            op = r[isrc].strip().lstrip("@!P0123456789 ").split(" ")[0].split(".")[0]
            ops[op] += n
        if total:
            f.write(f"\nWarp instructions executed: {total} ; average predicated-on lanes per instruction: {lanes / total:.1f}\n\n"
                    "| SASS opcode | warp instructions | share |\n|---|---:|---:|\n")
            for op, n in ops.most_common(18):
                f.write(f"| {op} | {n} | {n / total * 100:.1f}% |\n")
    print("wrote", kernel)

