"""Step time of a scene whose bodies all move (two-galaxy, BASELINE config 4's kind): what the mover count of
sort_repair.cuh costs where the repair never runs.   python scripts/moving_scene_step.py [n]   (NB_B200_NO_REPAIR=1 to compare)"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nbody_simulation_b200 as nb
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
w = nb.World(n, theta=0.5, dt=0.1)
w.generate_on_device(nb.NB_SCENE_TWO_GALAXY, 1, n)
w.set_counting(False)
w.step(3)
for _ in range(3):
    w.step(1)
best = None
for _ in range(3):
    c = nb.Counting()
    for _ in range(10):
        w.flush_l2()
        w.step(1, c)
    r = {"build_ms": c.build_bvh / 10 * 1e3, "traverse_ms": c.sum_gravity / 10 * 1e3}
    if best is None or r["build_ms"] < best["build_ms"]:
        best = r
w.build_tree()
ti = w.tree_info()
print(json.dumps({"n": n, "no_repair": os.environ.get("NB_B200_NO_REPAIR", "0"), **{k: round(v, 4) for k, v in best.items()},
                  "repaired": int(ti.sort_repaired), "passes": int(ti.sort_top_digits)}))
w.close()
