"""Small end-to-end case for compute-sanitizer (memcheck / racecheck): both trees, 3 steps, accel,
tree export, direct sum, draw.  Run under gpurun, one tool per call:
  compute-sanitizer --tool memcheck python scripts/sanitize_case.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nbody_simulation_b200 as nb

N = int(sys.argv[1]) if len(sys.argv) > 1 else 20_001
p = nb.generate_scene(nb.NB_SCENE_TWO_GALAXY, 1, N)
for tree in (nb.NB_TREE_QUAD, nb.NB_TREE_LBVH, nb.NB_TREE_BVH_REF):
    for mode in (nb.NB_FORCE_EXACT, nb.NB_FORCE_FAST):
        w = nb.World(len(p), theta=0.5, tree=tree, force_mode=mode)
        w.set_particles(p)
        c = nb.Counting()
        w.step(3, c)
        ax, ay = w.accel()
        t = w.tree()
        dx, dy, _ = w.accel_direct(10, 1000)
        f = w.draw()
        q = w.particles
        print(tree, mode, int(c.interactions), float(ax.sum()), int(t["info"].n_nodes), int(f.sum()))
        w.close()
w = nb.World(30_000, fast_walk_group=True)  # round 2: the group walk, the device generators, the statistics calls
w.generate_on_device(nb.NB_SCENE_PLUMMER, 3, 30_000)
w.step(2)
w.generate_on_device(nb.NB_SCENE_TWO_GALAXY, 3, 30_000)
w.step(1)
w.build_tree()
print(w.empty(), w.prune(), len(w.profile_step()))
w.close()
w = nb.World(200_000)
w.generate_on_device(nb.NB_SCENE_DEFAULT, 3, 0)
w.step(1)
w.close()
print("sanitize case done")
