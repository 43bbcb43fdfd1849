# N-GPU pass: tests/multi_gpu_check.py (bit-identity with one GPU, both exchanges, sharded call, fused walk + push,
# a failed depth guess inside a batch), then the bench line with and without the fused kernel / second stream.
#   gpurun --gpus N -- 'bash scripts/ab_multi.sh N'
cd $GRAFT_REPO_ROOT
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29511 tests/multi_gpu_check.py > gpurun_out/multi_gpu_check_$N.log 2>&1; echo "check rc=$?"
grep -v "^rank [1-9]" gpurun_out/multi_gpu_check_$N.log | tail -8
for nf in 1 0; do
NB_B200_NO_FUSE=$nf timeout 600 $TR --master-port 2951$((3+nf)) bench.py --gpus $N --steps 50 --warmup 5 > gpurun_out/bench_n${N}_nofuse$nf.json 2> gpurun_out/bench_n${N}_nofuse$nf.err; echo "bench nofuse=$nf rc=$?"; tail -2 gpurun_out/bench_n${N}_nofuse$nf.err
python -c "
import json
d = [json.loads(l) for l in open('gpurun_out/bench_n${N}_nofuse$nf.json') if l.startswith('{')][-1]; print(d['n_gpus'], round(d['ms_per_step'], 3), {k: round(v, 3) for k, v in d['phases_ms'].items()}, '%.3e' % d['value'], d.get('parity'), d.get('e2e',{}).get('ms_per_step'))"
done
