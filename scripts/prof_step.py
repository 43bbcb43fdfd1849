#!/usr/bin/env python
"""Per-kernel device times of one step (nb_profile_step): python scripts/prof_step.py [n] [tree]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nbody_simulation_b200 as nb
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
tree = {"quad": nb.NB_TREE_QUAD, "lbvh": nb.NB_TREE_LBVH}[sys.argv[2] if len(sys.argv) > 2 else "quad"]
w = nb.World(n, theta=0.5, tree=tree)
w.generate_on_device(nb.NB_SCENE_UNIFORM_DISC, 1, n)
w.set_counting(False)
w.step(5)
best = None
for _ in range(5):
    w.flush_l2()
    t = w.profile_step()
    if best is None or sum(u for _, u in t) < sum(u for _, u in best):
        best = t
w.build_tree(); ti = w.tree_info(); print("repaired", ti.sort_repaired, "movers", ti.sort_movers, "n_in", ti.n_in_tree, end=" | ")
print(n, " ".join(f"{k.split('_kernel')[0]}={u:.1f}" for k, u in best), "| build", round(sum(u for k, u in best if not k.startswith(("traverse", "integrate"))), 1))
w.close()
