# gpurun helper (round 2): every kernel of the step with its device time AND its DRAM bytes (one ncu pass, three
# metrics), for the 1M-body bench workload or any other bench arguments:
#   gpurun -- 'bash scripts/r02_ncu_kernels.sh r02_1m'        gpurun -- 'bash scripts/r02_ncu_kernels.sh r02_8m --n-per-gpu 8000000'
cd $GRAFT_REPO_ROOT
TAG=$1; shift
ARGS="--steps 2 --warmup 3 --no-cpu-baseline --no-e2e $@"
python bench.py $ARGS > gpurun_out/plain_$TAG.log 2>&1 && ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 4000 --csv --log-file gpurun_out/kernels_$TAG.csv python bench.py $ARGS > gpurun_out/ncu_kernels_$TAG.log 2>&1
echo rc=$?; tail -2 gpurun_out/ncu_kernels_$TAG.log | cut -c1-300
