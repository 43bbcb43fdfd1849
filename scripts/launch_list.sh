cd $GRAFT_REPO_ROOT
ARGS="--steps 3 --warmup 3 --no-cpu-baseline --no-e2e"
python bench.py $ARGS > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv python bench.py $ARGS > gpurun_out/ncu_list.log 2>&1
echo rc=$?
