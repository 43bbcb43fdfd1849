cd "${GRAFT_REPO_ROOT:-.}"
TAG=${1:-r02h}
python scripts/ab_step.py --group 2 --reps 1 --steps 3 > gpurun_out/plain_ab_$TAG.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:traverse_hybrid_kernel -s 2 -c 1 -f \
    -o gpurun_out/prof_${TAG}_traverse_hybrid_kernel python scripts/ab_step.py --group 2 --reps 1 --steps 3 > gpurun_out/ncu_full_$TAG.log 2>&1
echo "full rc=$?"
