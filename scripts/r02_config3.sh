cd $GRAFT_REPO_ROOT
last() { grep '^{' | tail -1; }
python bench.py --config 3 --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | last > gpurun_out/r02_config3.jsonl
python bench.py --config 3 --tree quad --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | last >> gpurun_out/r02_config3.jsonl
python bench.py --config 1 --tree quad --theta 0.5 --steps 100 --warmup 5 --no-cpu-baseline 2>/dev/null | last > gpurun_out/r02_config1_quad05.jsonl
python - <<'PY'
import json
for f in ("gpurun_out/r02_config3.jsonl", "gpurun_out/r02_config1_quad05.jsonl"):
    for l in open(f):
        d = json.loads(l)
        print(d["config"]["workload"][:64], "| ms/step %.3f" % d["ms_per_step"], {k: round(v, 3) for k, v in d["phases_ms"].items()}, "%.3e" % d["value"], "frac %.3f" % d["roofline"]["frac"])
PY
