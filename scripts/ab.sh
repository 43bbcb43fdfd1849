#!/bin/bash
# gpurun helper: scripts/ab_step.py with the in-tree library and with nbody_simulation_b200/libnbody_b200_base.so
cd "${GRAFT_REPO_ROOT:-.}"
for cfg in "$@"; do
  for lib in "" "$PWD/nbody_simulation_b200/libnbody_b200_base.so"; do
    NB_B200_LIB=$lib timeout 120 python scripts/ab_step.py $cfg
  done
done
