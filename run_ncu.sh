cd $GRAFT_REPO_ROOT
python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:traverse_kernel -s 8 -c 1 -o gpurun_out/prof_traverse \
   python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ncu2.log 2>&1
echo "ncu full rc=$?"; tail -2 gpurun_out/ncu2.log
