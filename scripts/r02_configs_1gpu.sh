# BASELINE.json configs that run on one GPU (round 2): one JSON line each into gpurun_out/r02_config*.jsonl
#   gpurun -- 'bash scripts/r02_configs_1gpu.sh'
cd $GRAFT_REPO_ROOT
mkdir -p gpurun_out
last() { grep '^{' | tail -1; }
# config 1: the repo's default scene.  The reference runs it on the CPU (oracle port: 100 steps headless at THETA = 50 as
# shipped; theta = 0.5 is ~15 s per step on the reference's own tree, so fewer steps are run and the line says how many)
python bench.py --impl reference --config 1 --steps 100 --warmup 3 | last > gpurun_out/r02_config1.jsonl
python bench.py --impl reference --config 1 --theta 0.5 --steps 100 --warmup 3 | last >> gpurun_out/r02_config1.jsonl
python bench.py --impl reference --config 1 --tree quad --steps 100 --warmup 3 | last >> gpurun_out/r02_config1.jsonl
# ... and the same scene on the device, the reference's own tree (bit for bit) and the quadtree
python bench.py --config 1 --steps 100 --warmup 5 --no-cpu-baseline 2>/dev/null | last >> gpurun_out/r02_config1.jsonl
python bench.py --config 1 --theta 0.5 --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | last >> gpurun_out/r02_config1.jsonl
python bench.py --config 1 --tree quad --steps 100 --warmup 5 --no-cpu-baseline 2>/dev/null | last >> gpurun_out/r02_config1.jsonl
# config 3: 4M-body Plummer, bvh variant (LBVH), build / traverse split in phases_ms; and the quadtree for comparison
python bench.py --config 3 --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | last > gpurun_out/r02_config3.jsonl
python bench.py --config 3 --tree quad --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | last >> gpurun_out/r02_config3.jsonl
wc -l gpurun_out/r02_config1.jsonl gpurun_out/r02_config3.jsonl
python - <<'PY'
import json
for f in ("gpurun_out/r02_config1.jsonl", "gpurun_out/r02_config3.jsonl"):
    for l in open(f):
        d = json.loads(l)
        print(d.get("impl", "ours"), d["config"]["workload"][:70], "| ms/step %.3f" % d["ms_per_step"], "steps", d["steps"],
              {k: round(v, 3) for k, v in d.get("phases_ms", {}).items()}, "%.3e int/s" % d["value"])
PY
