#!/usr/bin/env python
"""ab_step.py — A/B a kernel change on one GPU without torch: times the step of the default bench workload
(1M-body disc, quadtree, theta 0.5, FAST) with the library named by NB_B200_LIB (default: the in-tree build),
and prints a checksum of the accelerations so that two builds can be compared bit for bit.

  python scripts/ab_step.py                         # in-tree libnbody_b200.so
  NB_B200_LIB=/path/to/other.so python scripts/ab_step.py
"""
import argparse
import hashlib
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import nbody_simulation_b200 as nb  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=1_000_000)
ap.add_argument("--scene", default="disc", choices=["disc", "plummer"])
ap.add_argument("--tree", default="quad", choices=["quad", "lbvh", "bvhref"])
ap.add_argument("--theta", type=float, default=0.5)
ap.add_argument("--steps", type=int, default=20)
ap.add_argument("--reps", type=int, default=3)
ap.add_argument("--group", type=int, default=0, help="NB_FORCE_FAST walk: 0 depth-first, 1 group walk, 2 hybrid (group far field + depth-first near field)")
args = ap.parse_args()
scene = {"disc": nb.NB_SCENE_UNIFORM_DISC, "plummer": nb.NB_SCENE_PLUMMER}[args.scene]
tree = {"quad": nb.NB_TREE_QUAD, "lbvh": nb.NB_TREE_LBVH, "bvhref": nb.NB_TREE_BVH_REF}[args.tree]
w = nb.World(args.n, theta=args.theta, dt=0.1, tree=tree, force_mode=nb.NB_FORCE_FAST, fast_walk_group=args.group)
w.generate_on_device(scene, 1, args.n)
cc = nb.Counting()
ax, ay = w.accel(cc)
digest = hashlib.sha256(ax.tobytes() + ay.tobytes()).hexdigest()[:16]
w.set_counting(False)
best = None
for _ in range(args.reps):
    w.generate_on_device(scene, 1, args.n)
    w.step(3)
    c = nb.Counting()
    for _ in range(args.steps):
        w.flush_l2()
        w.step(1, c)
    r = {"build_ms": c.build_bvh / args.steps * 1e3, "traverse_ms": c.sum_gravity / args.steps * 1e3,
         "integrate_ms": c.post_calculations / args.steps * 1e3}
    r["step_ms"] = sum(r.values())
    if best is None or r["traverse_ms"] < best["traverse_ms"]:
        best = r
x, y, vx, vy, _ = w.get_soa()
state = hashlib.sha256(x.tobytes() + y.tobytes() + vx.tobytes() + vy.tobytes()).hexdigest()[:16]
print(json.dumps({"lib": os.environ.get("NB_B200_LIB", "in-tree"), "workload": f"{args.n} {args.scene} {args.tree} theta={args.theta}" + (f" walk{args.group}" if args.group else ""),
                  "interactions": int(cc.interactions), "accel_sha": digest, "state_sha": state,
                  **{k: round(v, 4) for k, v in best.items()}}))
w.close()
