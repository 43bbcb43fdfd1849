#!/bin/bash
# gpurun helper (round 2): one full ncu capture of the timed traversal kernel (group walk) of the 1M-disc step
cd "${GRAFT_REPO_ROOT:-.}"
TAG=${1:-r02a}
python scripts/ab_step.py --reps 1 --steps 3 > gpurun_out/plain_ab_$TAG.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:traverse_group_kernel -s 2 -c 1 -f \
    -o gpurun_out/prof_${TAG}_traverse_group_kernel python scripts/ab_step.py --reps 1 --steps 3 > gpurun_out/ncu_full_$TAG.log 2>&1
echo "full rc=$?"
cat gpurun_out/plain_ab_$TAG.log
