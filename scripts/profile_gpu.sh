#!/bin/bash
# Run on the GPU box (under gpurun): plain bench, then the ncu launch list, then one ncu --set full
# capture of the kernel named in $1 (default traverse_kernel).  Outputs land in gpurun_out/.
#   gpurun --timeout 900 -- 'bash scripts/profile_gpu.sh traverse_kernel'
cd "${GRAFT_REPO_ROOT:-.}"
K=${1:-traverse_kernel}
ARGS="--steps 3 --warmup 3 --no-cpu-baseline --no-e2e"
python bench.py $ARGS > gpurun_out/plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv \
    python bench.py $ARGS > gpurun_out/ncu_list.log 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:$K -s 8 -c 1 -o gpurun_out/prof_$K \
    python bench.py $ARGS > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
