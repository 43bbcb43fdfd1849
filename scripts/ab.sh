#!/bin/bash
# gpurun helper: scripts/ab_step.py with the in-tree library and with nbody_simulation_b200/libnbody_b200_base.so
# (a build of the code to compare against, e.g. `git stash; make -C nbody_simulation_b200/csrc; cp ...; git stash pop; make ...`;
# *.so files are git-ignored but travel to the GPU box)
cd "${GRAFT_REPO_ROOT:-.}"
for cfg in "$@"; do
  for lib in "" "$PWD/nbody_simulation_b200/libnbody_b200_base.so"; do
    [ -n "$lib" ] && [ ! -f "$lib" ] && { echo "no baseline build at $lib: skipped"; continue; }
    NB_B200_LIB=$lib timeout 120 python scripts/ab_step.py $cfg
  done
done
