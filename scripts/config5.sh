#!/bin/bash
# gpurun helper: BASELINE.json configs[4] on ONE B200 (the 8-GPU form needs 8x the box time):
#   1. theta sweep 0.3 .. 0.8 on a 64M-body Plummer scene: throughput, phase split, force error vs direct sum
#   2. per-kernel time + DRAM bytes of one 64M-body step (ncu, three metrics, --clock-control none)
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
N=${N:-64000000}
timeout 200 python scripts/theta_sweep.py --n $N > gpurun_out/sweep_$N.jsonl 2> gpurun_out/sweep_$N.log; echo "sweep rc=$?"
tail -12 gpurun_out/sweep_$N.log
timeout 150 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 300 --csv \
  --log-file gpurun_out/sweep_${N}_kernels.csv python scripts/theta_sweep.py --n $N --thetas 0.5 --steps 1 --warmup 0 --sample 256 \
  > gpurun_out/sweep_${N}_ncu.log 2>&1; echo "ncu rc=$?"
tail -3 gpurun_out/sweep_${N}_ncu.log
