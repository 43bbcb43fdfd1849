#!/usr/bin/env python
"""Design numbers of the group walk (csrc/traverse_group.cuh) from the CPU oracle's tree — no GPU needed:
  python scripts/group_walk_stats.py [scene] [n] [oracle tree kind 0=quad 1=lbvh] [theta]
Per group of G tree-order-adjacent bodies: nodes classified by the group's box, mixed nodes, items every body takes
("full") and items only some take ("partial"), the lane density of the partial items, and what a per-lane list
would achieve (mean / max items per body)."""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_binding as ob  # noqa: E402

scene = sys.argv[1] if len(sys.argv) > 1 else "disc"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1_000_000
kind = int(sys.argv[3]) if len(sys.argv) > 3 else 0
theta = float(sys.argv[4]) if len(sys.argv) > 4 else 0.5
parts = ob.generate_scene({"default": 0, "disc": 1, "plummer": 2, "galaxy": 3}[scene], 1, n)
t = ob.Tree(kind, parts["x"], parts["y"], parts["weight"])
L = ob.lib()
L.orc_group_bfs_stats.argtypes = [C.c_void_p, C.c_float, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double)]
names = ("groups cls acc_full acc_part all_open mixed full_items part_items part_pairs max_part batches pairs "
         "leaf_entries max_total levels").split()
np.set_printoptions(linewidth=200, precision=1, suppress=True)
for G in (32, 16, 8, 4):
    out, hist = (C.c_double * 15)(), (C.c_double * 66)()
    L.orc_group_bfs_stats(t.h, theta, G, out, hist)
    o = list(out)
    g = o[0]
    print(f"G = {G}: per group " + ", ".join(f"{k} {v / g:.1f}" for k, v in zip(names[1:], o[1:])))
    print(f"   pairs per body {o[11] / g / G:.1f}; full items carry {o[6] * G / o[11]:.3f} of the pairs; lane density of the "
          f"partial items {o[8] / max(o[7], 1) / G:.3f}; per-lane lists: mean / max partial {(o[8] / G) / max(o[9], 1):.3f}, "
          f"all items {(o[11] / G) / max(o[13], 1):.3f}")
    if G == 32:
        h = np.array(list(hist)).reshape(2, 33) / g
        print("   node items by popcount 1..32:", h[0][1:])
        print("   body items by popcount 1..32:", h[1][1:])
