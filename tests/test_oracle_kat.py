"""Known-answer tests that pin the CPU oracle to the reference's source lines (hand-derived from
/root/reference, SURVEY.md §4 — the reference itself ships no tests or fixtures)."""
import ctypes as C

import numpy as np

import oracle_binding as ob

f32 = np.float32


def pair(p1, p2, m, a0=(0.0, 0.0)):
    L = ob.lib()
    ax, ay = C.c_float(a0[0]), C.c_float(a0[1])
    L.orc_pair_force(p1[0], p1[1], p2[0], p2[1], m, C.byref(ax), C.byref(ay))
    return f32(ax.value), f32(ay.value)


def test_pair_force_l1_law():
    # src/main.rs:236-252: diff=(3,4), sum=7, d=25 -> accel += (30,40)/(7*25)
    ax, ay = pair((0, 0), (3, 4), 10.0)
    assert ax == f32(30.0) / f32(175.0) and ay == f32(40.0) / f32(175.0)


def test_pair_force_r2_clamp():
    # main.rs:245-249: d = 1e-4 -> clamped to 0.001; accel.x += 0.01*1/(0.01*0.001)
    ax, ay = pair((0, 0), (0.01, 0.0), 1.0)
    d = f32(0.01)
    expect = (d * f32(1.0)) / (d * f32(0.001))
    assert ax == expect and ay == 0.0
    assert abs(float(ax) - 1000.0) < 1e-3


def test_pair_force_skips_self_and_non_normal():
    # main.rs:238-243: !sum.is_normal() -> accel unchanged
    assert pair((5, 5), (5, 5), 9.0, (1.5, -2.5)) == (f32(1.5), f32(-2.5))
    assert pair((0, 0), (np.inf, 0), 9.0, (1.5, -2.5)) == (f32(1.5), f32(-2.5))
    assert pair((0, 0), (np.nan, 0), 9.0, (1.5, -2.5)) == (f32(1.5), f32(-2.5))
    assert pair((0, 0), (1e-40, 0), 9.0, (1.5, -2.5)) == (f32(1.5), f32(-2.5))  # subnormal sum


def test_integrate_kick_drift():
    # main.rs:420-422: v=(1,2), a=(10,20), dt=0.1 -> v=(2,4), x += (0.2,0.4)
    L = ob.lib()
    x, y = np.array([7.0], f32), np.array([-3.0], f32)
    vx, vy = np.array([1.0], f32), np.array([2.0], f32)
    ax, ay = np.array([10.0], f32), np.array([20.0], f32)
    L.orc_integrate(x.ctypes.data, y.ctypes.data, vx.ctypes.data, vy.ctypes.data, ax.ctypes.data, ay.ctypes.data, 1,
                    0.1)
    dt = f32(0.1)
    assert vx[0] == f32(1.0) + f32(10.0) * dt and vy[0] == f32(2.0) + f32(20.0) * dt
    assert x[0] == f32(7.0) + vx[0] * dt and y[0] == f32(-3.0) + vy[0] * dt


def test_opening_test():
    # main.rs:370-372 with bvh_tree.rs Rectangle{offset,size}: s^2 = max(w,h)^2 = 100
    L = ob.lib()
    acc = lambda px, py, th: L.orc_opening_accept(px, py, 0.0, 0.0, 10.0, 4.0, 5.0, 2.0, th)
    assert acc(30.0, 2.0, 0.5) == 1   # d=25: 100 < 625*0.25
    assert acc(24.9, 2.0, 0.5) == 0   # d=19.9: 100 < 99.0 false
    assert acc(5.0, 2.5, 50.0) == 0   # strictly inside the box -> always open
    assert acc(10.0, 2.0, 50.0) == 1  # on the edge: contains() is strict, so outside
    assert acc(5.0, 3.0, 50.0) == 0


def test_morton_child_index_and_midline_rule():
    # quad_tree.rs:172-179: child = (y>mid)<<1 | (x>mid); equality goes to the LOWER child
    L = ob.lib()
    key = lambda x, y: L.orc_morton_key(x, y, 0.0, 0.0, 100000.0)
    top = lambda k: k >> 46
    assert top(key(10.0, 10.0)) == 0 and top(key(60000.0, 10.0)) == 1
    assert top(key(10.0, 60000.0)) == 2 and top(key(60000.0, 60000.0)) == 3
    assert top(key(50000.0, 50000.0)) == 0
    assert top(key(np.nextafter(f32(50000.0), f32(1e9)), 50000.0)) == 1
    assert key(-1.0, 5.0) == 1 << 48 and key(5.0, 100000.0) == 1 << 48  # within_bounds, main.rs:223-226
    assert key(0.0, 0.0) == 0


def test_bvh_axis_choice_and_leaf_rules():
    # bvh_tree.rs:70-77 on 130 bodies laid along x: mean splits them evenly in x (hori=0), not in y
    n = 130
    x = np.arange(n, dtype=f32)
    y = np.zeros(n, f32)
    w = np.ones(n, np.uint32)
    t = ob.Tree(ob.KIND_LITERAL_BVH, x, y, w).export()
    assert t["node_leaf"][0] == 0 and t["node_count"][0] == n
    kids = [1, int(t["node_skip"][1])]
    assert sorted(int(t["node_count"][k]) for k in kids) == [65, 65]
    # left = predicate-true side = x > mean (bvh_tree.rs:74): first child holds the large x
    first = t["order"][int(t["node_start"][1])]
    assert x[first] > 64.5
    assert t["node_y0"][0] == 0.0 and t["node_y1"][0] == 0.0  # max seeded 0, size 0


def test_bvh_leaf_cog_is_unweighted_and_bbox_max_seed():
    # bvh_tree.rs:100-108,120-123: leaf {(0,0) w=1, (10,0) w=99} -> cog (5,0), mass 100
    x, y = np.array([0.0, 10.0], f32), np.array([0.0, 0.0], f32)
    w = np.array([1, 99], np.uint32)
    t = ob.Tree(ob.KIND_LITERAL_BVH, x, y, w).export()
    assert t["node_mass"][0] == 100
    leaves = np.nonzero(t["node_leaf"])[0]
    assert len(leaves) == 2  # always a Root on top (bvh_tree.rs:90), two leaves below
    # all-negative coordinates: max stays (0,0) (bvh_tree.rs:42,59)
    t = ob.Tree(ob.KIND_LITERAL_BVH, -x - 1, y - 1, w).export()
    assert t["node_x0"][0] == -11.0 and t["node_x1"][0] == f32(-11.0) + (f32(0.0) - f32(-11.0))


def test_quadtree_split_rule():
    # quad_tree.rs:159-165: the 9th body into a full leaf subdivides; cell internal <=> > 8 bodies
    rng = np.random.default_rng(1)
    x = rng.uniform(1, 40000, 8).astype(f32)
    y = rng.uniform(1, 40000, 8).astype(f32)
    t = ob.Tree(ob.KIND_LITERAL_QUAD, x, y, np.ones(8, np.uint32)).export()
    assert t["info"].n_nodes == 1 and t["node_leaf"][0] == 1
    x = np.append(x, f32(90000.0))
    y = np.append(y, f32(90000.0))
    t = ob.Tree(ob.KIND_LITERAL_QUAD, x, y, np.ones(9, np.uint32)).export()
    assert t["node_leaf"][0] == 0 and t["node_count"][0] == 9
    assert (t["node_count"][t["node_leaf"] == 1] <= 8).all()


def test_u32_mass_wraps():
    # bvh_tree.rs:148 / quad_tree.rs:253-257: release-mode u32 adds wrap
    x = np.linspace(10, 90000, 20).astype(f32)
    w = np.full(20, 3_000_000_000, np.uint32)
    t = ob.Tree(ob.KIND_LITERAL_QUAD, x, x, w).export()
    assert int(t["node_mass"][0]) == (20 * 3_000_000_000) % (1 << 32)


def test_quad_empty_and_prune_counts():
    # quad_tree.rs:66-137 on a fresh tree: empty() visits every node once; then every child of
    # every visited Root is empty, so prune() removes the root's children only
    p_x = np.random.default_rng(2).uniform(0, 99999, 500).astype(f32)
    p_y = np.random.default_rng(3).uniform(0, 99999, 500).astype(f32)
    T = ob.Tree(ob.KIND_LITERAL_QUAD, p_x, p_y, np.ones(500, np.uint32))
    info = T.info()
    emptied, pruned = T.empty_prune()
    assert emptied == info.n_nodes
    assert 1 <= pruned <= 4
