#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/ab.log
for n in 1000000 2000000 4000000 8000000; do
  for lib in /root/repo/ab_old.so ""; do
    NB_B200_LIB=$lib python scripts/ab_step.py --n $n --reps 3 --steps 20 >> gpurun_out/ab.log 2>&1
  done
done
python -m pytest tests/test_gpu_parity.py -x -q -m gpu > gpurun_out/ab_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/ab.log
cut -c1-40,150- gpurun_out/ab.log; tail -3 gpurun_out/ab_tests.log
