#!/bin/bash
# gpurun helper (end of round 1): launch list of the bench command + one full capture of the timed traversal kernel
cd "${GRAFT_REPO_ROOT:-.}"
bash scripts/launch_list.sh r01e
# ab_step.py launches traverse_fast_kernel<1> once (counted accel), then the timed <0> variant: skip 2, take 1
python scripts/ab_step.py --reps 1 --steps 3 > gpurun_out/plain_ab_r01e.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:traverse_fast_kernel -s 2 -c 1 -f \
    -o gpurun_out/prof_r01e_traverse_fast_kernel python scripts/ab_step.py --reps 1 --steps 3 > gpurun_out/ncu_full_r01e.log 2>&1
echo "full rc=$?"
tail -2 gpurun_out/plain_r01e.log
