# gpurun helper: the committed ncu evidence at the end-of-round-2 code state
cd $GRAFT_REPO_ROOT
TAG=${1:-r02g}
bash scripts/r02_ncu_kernels.sh ${TAG}_1m
python scripts/ab_step.py --reps 1 --steps 3 > gpurun_out/plain_ab_$TAG.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:traverse_fast_kernel -s 2 -c 1 -f \
    -o gpurun_out/prof_${TAG}_traverse_fast_kernel python scripts/ab_step.py --reps 1 --steps 3 > gpurun_out/ncu_full_$TAG.log 2>&1
echo "full rc=$?"
python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/ref_$TAG.json 2> gpurun_out/ref_$TAG.err; echo ref rc=$?
python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo bench rc=$?
