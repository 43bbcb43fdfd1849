#!/bin/bash
# gpurun helper: GPU parity tests, then a short bench (prints the key numbers).
cd "${GRAFT_REPO_ROOT:-.}"
timeout 600 python -m pytest tests -q -m gpu 2>&1 | tail -5
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
python -c "
import json; d=json.load(open('gpurun_out/bench.json')); print({k:d[k] for k in ('value','ms_per_step','phases_ms','steps_per_s')}); print('frac', d['roofline']['frac'], 'e2e ms', d.get('e2e',{}).get('ms_per_step'))"
