#!/bin/bash
# gpurun helper: the end-of-round measurement set on one B200
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 120 python scripts/theta_sweep.py --n 64000000 > gpurun_out/sweep_64000000.jsonl 2> gpurun_out/sweep_64000000.log; echo "sweep rc=$?"
timeout 200 python bench.py > gpurun_out/bench_r01e.json 2> gpurun_out/bench_r01e.err; echo "bench rc=$?"
bash scripts/profile_r01e.sh
