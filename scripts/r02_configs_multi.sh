# BASELINE.json configs 4 and 5 on N GPUs of one box (round 2): one JSON line per run, appended to gpurun_out/
#   gpurun --gpus N -- 'bash scripts/r02_configs_multi.sh N [4|5|45]'
cd $GRAFT_REPO_ROOT
N=${1:-2}
WHICH=${2:-4}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
last() { grep '^{' | tail -1; }
if [[ $WHICH == *4* ]]; then   # 16M two-galaxy collision, Morton-range sharded, strong-scaled over N
  timeout 900 $TR --master-port 29521 bench.py --gpus $N --config 4 --steps 10 --warmup 3 2>gpurun_out/r02_config4_n$N.err | last >> gpurun_out/r02_config4.jsonl
fi
if [[ $WHICH == *5* ]]; then   # 64M Plummer theta sweep
  for th in 0.3 0.5 0.8; do
    timeout 900 $TR --master-port 29522 bench.py --gpus $N --config 5 --theta $th --steps 5 --warmup 3 --no-e2e 2>gpurun_out/r02_config5_n$N.err | last >> gpurun_out/r02_config5.jsonl
  done
fi
python - <<'PY'
import json, glob
for f in sorted(glob.glob("gpurun_out/r02_config[45].jsonl")):
    for l in open(f):
        d = json.loads(l)
        print(d["n_gpus"], d["config"]["workload"][:60], "theta", d["config"]["theta"], "| ms/step %.3f" % d["ms_per_step"],
              {k: round(v, 3) for k, v in d.get("phases_ms", {}).items()}, "%.3e int/s" % d["value"], d.get("parity", {}).get("bit_identical_to_1gpu"),
              "e2e ms", d.get("e2e", {}).get("ms_per_step"))
PY
