# gpurun helper: ncu launch list of a bench configuration ->  gpurun_out/launches_<tag>.csv
#   gpurun -- 'bash scripts/launch_list.sh bvhref --tree bvhref'
cd $GRAFT_REPO_ROOT
TAG=$1; shift
ARGS="--steps 2 --warmup 3 --no-cpu-baseline --no-e2e $@"
python bench.py $ARGS > gpurun_out/plain_$TAG.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/launches_$TAG.csv python bench.py $ARGS > gpurun_out/ncu_list_$TAG.log 2>&1
echo rc=$?
