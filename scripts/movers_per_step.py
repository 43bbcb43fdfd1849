"""How many bodies leave their leaf cell per step (sort_repair.cuh) and whether the builds repair or sort:
  python scripts/movers_per_step.py 8000000 disc"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nbody_simulation_b200 as nb
n = int(sys.argv[1]); scene = {"disc": nb.NB_SCENE_UNIFORM_DISC, "plummer": nb.NB_SCENE_PLUMMER}[sys.argv[2] if len(sys.argv) > 2 else "disc"]
w = nb.World(n, theta=0.5)
w.generate_on_device(scene, 1, n)
w.set_counting(False)
out = []
for s in range(14):
    w.step(1)
    w.build_tree()
    ti = w.tree_info()
    out.append((ti.sort_repaired, ti.sort_movers, ti.sort_redone, ti.n_in_tree == n))
print(n, sys.argv[2:] , out)
w.step(200)
w.build_tree(); ti = w.tree_info(); print("after 200 more:", ti.sort_repaired, ti.sort_movers, ti.sort_redone)
w.close()
