#!/bin/bash
# A/B: old vs new library on build phases; then parity tests touching the sort / keys
mkdir -p gpurun_out
for n in 1000000 8000000; do
  for lib in /root/repo/ab_old.so ""; do
    NB_B200_LIB=$lib python scripts/ab_step.py --n $n --reps 3 --steps 20 >> gpurun_out/ab_morton.log 2>&1
  done
done
python -m pytest tests/test_gpu_parity.py -x -q -m gpu > gpurun_out/ab_morton_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/ab_morton.log
cat gpurun_out/ab_morton.log; tail -3 gpurun_out/ab_morton_tests.log
