# 2-GPU pass: parity at world size 2 (both exchanges), then the weak-scaled bench line with each exchange.
#   gpurun --gpus 2 -- 'bash scripts/multi2.sh'
cd $GRAFT_REPO_ROOT
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $TR --master-port 29511 tests/multi_gpu_check.py 2>&1 | grep -v "^rank [1-9]" | tail -16
for ex in nccl peer; do
  timeout 300 $TR --master-port 29512 bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline --no-e2e --exchange $ex 2>/dev/null | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read()); print('$ex', d['n_gpus'], round(d['ms_per_step'], 3), {k: round(v, 3) for k, v in d['phases_ms'].items()}, '%.3e' % d['value'])"
done
