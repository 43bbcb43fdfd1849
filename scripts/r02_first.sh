#!/bin/bash
# gpurun helper (round 2): parity of the group walk, then A/B against the depth-first walk
cd "${GRAFT_REPO_ROOT:-.}"
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_golden.py -x -q -m gpu 2>&1 | tail -15
for a in "" "--dfs"; do timeout 300 python scripts/ab_step.py $a; done
timeout 300 python scripts/ab_step.py --tree lbvh
timeout 300 python scripts/ab_step.py --tree lbvh --dfs
timeout 300 python scripts/ab_step.py --scene plummer --n 4000000
timeout 300 python scripts/ab_step.py --scene plummer --n 4000000 --dfs
