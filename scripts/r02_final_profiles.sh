# gpurun helper: the committed ncu evidence at the end-of-round-2 code state
cd $GRAFT_REPO_ROOT
bash scripts/r02_ncu_kernels.sh r02f_1m
bash scripts/r02_ncu_kernels.sh r02f_8m --n-per-gpu 8000000
python scripts/ab_step.py --reps 1 --steps 3 > gpurun_out/plain_ab_r02f.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:traverse_fast_kernel -s 2 -c 1 -f \
    -o gpurun_out/prof_r02f_traverse_fast_kernel python scripts/ab_step.py --reps 1 --steps 3 > gpurun_out/ncu_full_r02f.log 2>&1
echo "full rc=$?"
