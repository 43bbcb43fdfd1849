# 8-GPU pass: parity check at world size 8 (ncclAllGather and peer-store exchange), the default weak-scaled bench
# line with each exchange, and BASELINE config 4 (16M-body two-galaxy scene, 2M bodies per GPU).
#   gpurun --gpus 8 -- 'bash scripts/multi8.sh'
cd $GRAFT_REPO_ROOT
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
mkdir -p gpurun_out
timeout 300 $TR --master-port 29511 tests/multi_gpu_check.py 2>&1 | tail -3
timeout 300 $TR --master-port 29512 bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | tail -1 | tee gpurun_out/multi${N}_default.json
echo
timeout 300 $TR --master-port 29514 bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline --no-e2e --exchange nccl 2>/dev/null | tail -1 | tee gpurun_out/multi${N}_default_nccl.json
echo
timeout 400 $TR --master-port 29513 bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --scene galaxy --n-per-gpu $((16000000 / N)) 2>/dev/null | tail -1 | tee gpurun_out/multi${N}_galaxy16m.json
