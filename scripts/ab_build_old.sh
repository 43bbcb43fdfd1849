#!/bin/bash
# builds the library of git HEAD into /root/repo/ab_old.so (A/B partner of scripts/ab_step.py)
set -e
rm -rf /tmp/oldsrc && mkdir -p /tmp/oldsrc/nbody_simulation_b200/csrc /tmp/oldsrc/include
cd /root/repo
for f in $(git ls-files nbody_simulation_b200/csrc include); do mkdir -p /tmp/oldsrc/$(dirname $f); git show HEAD:$f > /tmp/oldsrc/$f; done
make -C /tmp/oldsrc/nbody_simulation_b200/csrc OUT=/root/repo/ab_old.so 2>&1 | tail -1
