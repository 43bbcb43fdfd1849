"""CPU tier: the oracle's two independent formulations agree with each other.
(A) literal quad_tree.rs (insert/subdivide/calculate_gravity) flattened in pre-order, and
(B) the Morton-key -> stable sort -> range-split pipeline the GPU implements.
If these agree bit for bit, the GPU (which is compared with B) is pinned to the literal
reference algorithm A."""
import numpy as np
import pytest

import nbody_simulation_b200 as nb
import oracle_binding as ob

KEYS = ["order", "node_depth", "node_start", "node_count", "node_skip", "node_parent", "node_leaf", "node_mass",
        "node_cogx", "node_cogy", "node_x0", "node_y0", "node_x1", "node_y1", "node_s2"]


def bits(a):
    return np.ascontiguousarray(a).view(np.uint8)


SCENES = [(nb.NB_SCENE_UNIFORM_DISC, 60_000), (nb.NB_SCENE_DEFAULT, 0), (nb.NB_SCENE_PLUMMER, 80_000),
          (nb.NB_SCENE_TWO_GALAXY, 50_001)]


@pytest.mark.parametrize("scene,n", SCENES)
def test_literal_quadtree_equals_morton_pipeline(scene, n):
    p = nb.generate_scene(scene, 11, n)
    a = ob.Tree(ob.KIND_LITERAL_QUAD, p["x"], p["y"], p["weight"]).export()
    b = ob.Tree(ob.KIND_PIPE_QUAD, p["x"], p["y"], p["weight"]).export()
    assert a["info"].n_nodes == b["info"].n_nodes and a["info"].n_in_tree == b["info"].n_in_tree
    for k in KEYS:
        assert np.array_equal(bits(a[k]), bits(b[k])), k


def test_literal_recursive_walk_equals_flat_stackless_walk():
    p = nb.generate_scene(nb.NB_SCENE_DEFAULT, 3)[:30_000]
    for kind in (ob.KIND_LITERAL_QUAD, ob.KIND_LITERAL_BVH):
        T = ob.Tree(kind, p["x"], p["y"], p["weight"])
        for theta in (0.5, 50.0):
            ax, ay, ni, nt, _ = T.accel(theta, literal=True)
            bx, by, mi, mt, _ = T.accel(theta, literal=False)
            assert (ni, nt) == (mi, mt)
            assert np.array_equal(bits(ax), bits(bx)) and np.array_equal(bits(ay), bits(by))


def test_results_do_not_depend_on_thread_count():
    # rayon's ordered collect (main.rs:406-416): per-body walks are sequential
    p = nb.generate_scene(nb.NB_SCENE_UNIFORM_DISC, 4, 20_000)
    T = ob.Tree(ob.KIND_PIPE_QUAD, p["x"], p["y"], p["weight"])
    a = T.accel(0.5, threads=1)
    b = T.accel(0.5, threads=4)
    assert a[2:4] == b[2:4] and np.array_equal(bits(a[0]), bits(b[0])) and np.array_equal(bits(a[1]), bits(b[1]))


@pytest.mark.parametrize("kind", [ob.KIND_PIPE_QUAD, ob.KIND_PIPE_LBVH])
def test_tree_invariants(kind):
    p = nb.generate_scene(nb.NB_SCENE_PLUMMER, 5, 40_000)
    t = ob.Tree(kind, p["x"], p["y"], p["weight"]).export()
    nn = int(t["info"].n_nodes)
    cap = 8 if kind == ob.KIND_PIPE_QUAD else 64
    leaf = t["node_leaf"] == 1
    assert sorted(t["order"].tolist()) == list(range(len(p)))          # a permutation
    assert t["node_count"][leaf].sum() == t["info"].n_in_tree           # leaves tile the bodies
    assert (t["node_count"][leaf] <= cap).all() and (t["node_count"][~leaf] > cap).all()
    assert t["node_skip"][0] == nn and t["node_parent"][0] == 0xFFFFFFFF
    assert (t["node_skip"][leaf] == np.arange(nn)[leaf] + 1).all()     # pre-order
    assert t["node_mass"][0] == p["weight"].sum() % (1 << 32)
    par = t["node_parent"][1:]
    assert (par < np.arange(1, nn)).all()
    # children partition the parent's range, in order
    kids = {}
    for i in range(1, nn):
        kids.setdefault(int(par[i - 1]), []).append(i)
    for q, ch in list(kids.items())[:2000]:
        assert t["node_start"][ch[0]] == t["node_start"][q]
        assert sum(int(t["node_count"][c]) for c in ch) == t["node_count"][q]
        assert len(ch) <= (4 if kind == ob.KIND_PIPE_QUAD else 2)
        if kind == ob.KIND_PIPE_LBVH:
            assert len(ch) == 2  # path-compressed radix tree: every internal node branches


def test_fp32_tree_force_against_fp64_direct_sum():
    # Barnes-Hut error level of the reference criterion at theta = 0.5 (documentation, not parity)
    p = nb.generate_scene(nb.NB_SCENE_UNIFORM_DISC, 6, 30_000)
    tg = np.arange(0, 30_000, 15, dtype=np.uint32)
    rx, ry = ob.direct_f64(p["x"], p["y"], p["weight"], tg)
    for kind in (ob.KIND_PIPE_QUAD, ob.KIND_PIPE_LBVH, ob.KIND_LITERAL_BVH):
        ax, ay, _, _, _ = ob.Tree(kind, p["x"], p["y"], p["weight"]).accel(0.5, tg)
        assert ob.normwise_err(ax, ay, rx, ry) < 5e-3
    fx, fy = ob.direct_f32(p["x"], p["y"], p["weight"], tg)
    assert ob.normwise_err(fx, fy, rx, ry) < 1e-3  # fp32 sequential accumulation noise floor


def test_world_update_identity_vs_faithful_index_bug():
    # main.rs:419-423 applies acceleration[idx] to the PERMUTED particle idx.
    p = nb.generate_scene(nb.NB_SCENE_UNIFORM_DISC, 8, 5_000)
    a, b = ob.OracleWorld(p), ob.OracleWorld(p)
    a.update(ob.KIND_LITERAL_BVH, 0.1, 50.0, faithful_bug=False)
    b.update(ob.KIND_LITERAL_BVH, 0.1, 50.0, faithful_bug=True)
    xa, _, vxa, _, wa = a.get()
    xb, _, vxb, _, wb = b.get()
    assert np.array_equal(wa, wb)
    assert not np.array_equal(vxa, vxb)  # the bug is observable
    # identity semantics: every body got ITS OWN acceleration
    T = ob.Tree(ob.KIND_LITERAL_BVH, p["x"], p["y"], p["weight"])
    ax, _, _, _, _ = T.accel(50.0, literal=True)
    assert np.array_equal(bits(vxa), bits(np.float32(p["vx"] + ax * np.float32(0.1))))
