# The driver's N-GPU bench command, with a one-line summary.   gpurun --gpus N -- 'bash scripts/bench_multi.sh N'
cd $GRAFT_REPO_ROOT
N=${1:-8}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29512 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_n$N.err
python -c "
import json
d = [json.loads(l) for l in open('gpurun_out/bench_n$N.json') if l.startswith('{')][-1]; print(d['n_gpus'], round(d['ms_per_step'], 3), {k: round(v, 3) for k, v in d['phases_ms'].items()}, '%.3e' % d['value'], d.get('parity'), d.get('e2e'))"
