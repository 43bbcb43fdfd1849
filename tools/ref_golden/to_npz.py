#!/usr/bin/env python
"""Step 4 of tools/ref_golden: the raw dumps of the Rust test -> reference_golden.npz in the layout
tests/test_golden.py::test_oracle_matches_the_rust_reference reads (node arrays in pre-order like nb_get_tree)."""
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
meta = dict(l.split() for l in open(os.path.join(HERE, "ref_meta.txt")))
n, nn = int(meta["particles"]), int(meta["nodes"])
nodes = np.fromfile(os.path.join(HERE, "ref_nodes.bin"), dtype=np.dtype(
    [("depth", "<u4"), ("leaf", "<u4"), ("start", "<u4"), ("count", "<u4"), ("mass", "<u4"), ("cogx", "<f4"),
     ("cogy", "<f4"), ("ox", "<f4"), ("oy", "<f4"), ("sx", "<f4"), ("sy", "<f4")]))
assert len(nodes) == nn
count = nodes["count"].astype(np.int64).copy()
skip = np.zeros(nn, np.uint32)
stack = []  # pre-order: a node's subtree ends where a node of depth <= its own begins
for i in range(nn):
    while stack and nodes["depth"][stack[-1]] >= nodes["depth"][i]:
        skip[stack.pop()] = i
    stack.append(i)
for i in stack:
    skip[i] = nn
for i in range(nn - 1, -1, -1):  # counts of the Roots = sum over the leaves below
    if not nodes["leaf"][i]:
        c, j = 0, i + 1
        while j < skip[i]:
            c += count[j]
            j = skip[j]
        count[i] = c
rec = np.dtype([("x", "<f4"), ("y", "<f4"), ("vx", "<f4"), ("vy", "<f4"), ("weight", "<u4")])
vec = np.fromfile(os.path.join(HERE, "ref_vec.bin"), dtype=rec)
traj = np.fromfile(os.path.join(HERE, "ref_traj.bin"), dtype=rec)
acc = np.fromfile(os.path.join(HERE, "ref_accel.bin"), dtype="<f4").reshape(n, 2)
np.savez_compressed(os.path.join(HERE, "reference_golden.npz"), theta=np.float32(meta["theta"]),
                    dt=np.float32(meta["dt"]), steps=np.int32(meta["steps"]), node_depth=nodes["depth"].astype(np.int32),
                    node_leaf=nodes["leaf"].astype(np.uint8), node_start=nodes["start"], node_count=count.astype(np.uint32),
                    node_skip=skip, node_mass=nodes["mass"], node_cogx=nodes["cogx"], node_cogy=nodes["cogy"],
                    node_x0=nodes["ox"], node_y0=nodes["oy"], node_sx=nodes["sx"], node_sy=nodes["sy"],
                    vec=vec.view(np.uint8), traj=traj.view(np.uint8), ax=acc[:, 0].copy(), ay=acc[:, 1].copy())
print("wrote reference_golden.npz:", n, "particles,", nn, "nodes, theta", meta["theta"])
