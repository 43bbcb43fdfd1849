#!/bin/bash
mkdir -p gpurun_out; rm -f gpurun_out/ab.log
python -m pytest tests/test_gpu_parity.py -x -q -m gpu > gpurun_out/ab_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/ab.log
for n in 1000000 2000000 4000000 8000000; do
  for lib in /root/repo/ab_old.so ""; do
    NB_B200_LIB=$lib python scripts/ab_step.py --n $n --reps 3 --steps 20 >> gpurun_out/ab.log 2>&1
  done
done
python - <<'PY'
import json
for l in open('gpurun_out/ab.log'):
    if l.startswith('{'):
        d=json.loads(l); print(d['workload'], d['lib'][-10:], d['accel_sha'], d['state_sha'], d['build_ms'], d['traverse_ms'])
    else: print(l.strip())
PY
tail -3 gpurun_out/ab_tests.log
