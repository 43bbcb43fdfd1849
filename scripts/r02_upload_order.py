#!/usr/bin/env python
"""What the first step after an upload costs (round 2): N bodies in CALLER order (a random
permutation of a tree-ordered state, like any caller's Vec), two nb_world_update calls; prints the step's phase times.
  NB_B200_AOS_PERMUTE_MIN=4000000000 python scripts/r02_upload_order.py 64000000     # round-1 behaviour
  python scripts/r02_upload_order.py 64000000                                          # re-order through the AoS copy"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import nbody_simulation_b200 as nb
n = int(sys.argv[1]) if len(sys.argv) > 1 else 16_000_000
w = nb.World(n, theta=0.5)
w.generate_on_device(nb.NB_SCENE_UNIFORM_DISC, 1, n)
p = w.particles                      # generation order: uncorrelated with position, like a caller's Vec
host, owner = nb.pinned_particles(n)
host[:] = p
w.set_counting(False)
for rep in range(3):
    c = nb.Counting()
    t0 = time.perf_counter()
    w.update_host(host, 0.1, c)
    wall = time.perf_counter() - t0
    print(f"n={n} call {rep}: wall {wall * 1e3:.1f} ms, build {c.build_bvh * 1e3:.2f} traverse {c.sum_gravity * 1e3:.2f} "
          f"integrate {c.post_calculations * 1e3:.2f} ms (permute from {os.environ.get('NB_B200_AOS_PERMUTE_MIN', 'default 2M')})")
w.close()
