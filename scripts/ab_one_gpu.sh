#!/bin/bash
# One-GPU pass used while developing a build-side change: parity tests, per-kernel times, movers per step.
# (A/B against the library of git HEAD: scripts/ab_build_old.sh builds it into ab_old.so, NB_B200_LIB selects it.)
mkdir -p gpurun_out; rm -f gpurun_out/ab.log
python -m pytest tests/test_gpu_parity.py -x -q -m gpu > gpurun_out/ab_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/ab.log
for n in 1000000 8000000; do
  python scripts/prof_step.py $n >> gpurun_out/ab.log 2>&1
  python scripts/_movers.py $n disc >> gpurun_out/ab.log 2>&1
done
python scripts/_movers.py 4000000 plummer >> gpurun_out/ab.log 2>&1
cat gpurun_out/ab.log
tail -3 gpurun_out/ab_tests.log
