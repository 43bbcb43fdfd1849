#!/bin/bash
# gpurun helper: one `ncu --set full` capture per named kernel of a bench configuration.
#   gpurun -- 'SKIP=7 bash scripts/profile_kernels.sh r01d "" traverse_fast_kernel'   (SKIP = matching launches to skip; default 4.
#   bench.py launches traverse_fast_kernel<1> (counters on) for its replay: 7 skips reach the timed <0> variant)
cd "${GRAFT_REPO_ROOT:-.}"
TAG=$1; EXTRA=$2; shift 2
ARGS="--steps 2 --warmup 3 --no-cpu-baseline --no-e2e $EXTRA"
python bench.py $ARGS > gpurun_out/plain_$TAG.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_$TAG.log; exit 1; }
for K in "$@"; do
  ncu --set full --clock-control none --import-source on -k regex:$K -s ${SKIP:-4} -c 1 -f -o gpurun_out/prof_${TAG}_$K \
      python bench.py $ARGS > gpurun_out/ncu_full_${TAG}_$K.log 2>&1
  echo "$K rc=$?"
done
