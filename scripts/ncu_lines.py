#!/usr/bin/env python
"""Per-source-line instruction counts of one kernel from an .ncu-rep (run here, no GPU needed):
  python scripts/ncu_lines.py gpurun_out/prof.ncu-rep <mangled-name-substring> [lib.so]
Joins ncu's per-SASS-instruction "Instructions Executed" with nvdisasm -g line info of the in-tree library (same build!)."""
import csv, os, re, subprocess, sys, tempfile, collections
rep, pat = sys.argv[1], sys.argv[2]
lib = sys.argv[3] if len(sys.argv) > 3 else os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "nbody_simulation_b200", "libnbody_b200.so")
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr = rows[1]
iS, iN, iT, iP = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("Predicated-On Thread Instructions Executed")
iSamp = hdr.index("# Samples")
prof = [(r[iS].strip(), int(r[iN]), int(r[iT]), int(r[iP]), int(r[iSamp])) for r in rows[2:] if len(r) > iP]
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", lib], cwd=tmp, capture_output=True)
cub = [f for f in os.listdir(tmp) if f.endswith(".cubin") and "scene" not in f][0]
dis = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cub)], capture_output=True, text=True).stdout.splitlines()
start = next(i for i, l in enumerate(dis) if l.startswith(".text.") and pat in l)
line, insts = None, []
for l in dis[start + 1:]:
    if l.startswith("//-----") or l.startswith("\t.section"): break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m: line = (os.path.basename(m.group(1)), int(m.group(2))); continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(.*?);", l)
    if m: insts.append((line, m.group(1).strip()))
assert len(insts) == len(prof), (len(insts), len(prof))
by = collections.OrderedDict()
tot = sum(p[1] for p in prof)
for (ln, txt), (src, n, t, pon, smp) in zip(insts, prof):
    d = by.setdefault(ln, [0, 0, 0, 0, 0])
    d[0] += n; d[1] += t; d[2] += pon; d[3] += smp; d[4] += 1
print(f"total warp instructions {tot}")
for ln, d in sorted(by.items(), key=lambda kv: (kv[0][0], kv[0][1])):
    print(f"{ln[0]}:{ln[1]:4d}  sass {d[4]:4d}  warp-inst {d[0]:12d} ({100*d[0]/tot:5.1f}%)  lanes-on {d[2]/max(d[0],1):5.1f}  samples {d[3]}")
