"""Committed golden vectors (tests/golden/golden_small.npz, made by tests/golden/make_golden.py from
the oracle): the oracle must keep reproducing them (CPU tier) and the CUDA path must reproduce them
bit for bit through the C ABI (GPU tier)."""
import os

import numpy as np
import pytest

import nbody_simulation_b200 as nb
import oracle_binding as ob

G = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "golden_small.npz"))
P = np.ascontiguousarray(G["particles"]).view(nb.PARTICLE_DTYPE)
TREE_KEYS = ["keys", "order", "node_depth", "node_start", "node_count", "node_skip", "node_parent", "node_leaf",
             "node_mass", "node_cogx", "node_cogy", "node_x0", "node_y0", "node_x1", "node_y1", "node_s2"]


def bits(a):
    return np.ascontiguousarray(a).view(np.uint8)


@pytest.mark.parametrize("name,kind", [("quad", ob.KIND_PIPE_QUAD), ("lbvh", ob.KIND_PIPE_LBVH),
                                       ("bvhref", ob.KIND_LITERAL_BVH)])
def test_oracle_reproduces_golden(name, kind):
    T = ob.Tree(kind, P["x"], P["y"], P["weight"])
    t = T.export()
    for k in TREE_KEYS:
        assert np.array_equal(bits(t[k]), bits(G[f"{name}_{k}"])), k
    ax, ay, ni, nt, _ = T.accel(float(G["theta"]), literal=(kind == ob.KIND_LITERAL_BVH))
    assert ni == int(G[f"{name}_interactions"]) and nt == int(G[f"{name}_tests"])
    assert np.array_equal(bits(ax), bits(G[f"{name}_ax"])) and np.array_equal(bits(ay), bits(G[f"{name}_ay"]))
    w = ob.OracleWorld(P)
    for _ in range(int(G["steps"])):
        w.update(kind, float(G["dt"]), float(G["theta"]))
    x, y, vx, vy, _ = w.get()
    assert np.array_equal(bits(np.stack([x, y, vx, vy])), bits(G[f"{name}_traj"]))


def test_golden_literal_quadtree_matches_too():
    lit = ob.Tree(ob.KIND_LITERAL_QUAD, P["x"], P["y"], P["weight"]).export()
    for k in TREE_KEYS[1:]:
        assert np.array_equal(bits(lit[k]), bits(G[f"quad_{k}"])), k


@pytest.mark.gpu
@pytest.mark.parametrize("name,tree", [("quad", nb.NB_TREE_QUAD), ("lbvh", nb.NB_TREE_LBVH),
                                       ("bvhref", nb.NB_TREE_BVH_REF)])
def test_cuda_reproduces_golden(name, tree):
    w = nb.World(len(P), theta=float(G["theta"]), dt=float(G["dt"]), tree=tree, force_mode=nb.NB_FORCE_EXACT)
    w.set_particles(P)
    c = nb.Counting()
    ax, ay = w.accel(c)
    t = w.tree()
    for k in TREE_KEYS[1:] if name == "bvhref" else TREE_KEYS:  # the reference BVH has no Morton keys
        assert np.array_equal(bits(t[k]), bits(G[f"{name}_{k}"])), k
    assert c.interactions == int(G[f"{name}_interactions"]) and c.node_tests == int(G[f"{name}_tests"])
    assert np.array_equal(bits(ax), bits(G[f"{name}_ax"])) and np.array_equal(bits(ay), bits(G[f"{name}_ay"]))
    w.step(int(G["steps"]))
    q = w.particles
    assert np.array_equal(bits(np.stack([q["x"], q["y"], q["vx"], q["vy"]])), bits(G[f"{name}_traj"]))
    w.close()
