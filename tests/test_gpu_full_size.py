"""GPU parity at the BASELINE.json sizes (1M quadtree, 4M Plummer LBVH, 16M two-galaxy quadtree).
The oracle builds the full tree on the CPU (seconds) and walks a SAMPLE of bodies; the CUDA path must
match bit for bit in EXACT mode (tree arrays in full, forces on the sample), and FAST must stay within
1e-5 (normwise) of EXACT over ALL bodies.  Size-independent invariants are checked on the full arrays."""
import numpy as np
import pytest

import nbody_simulation_b200 as nb
import oracle_binding as ob

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a).view(np.uint8)


CASES = [
    ("disc-1M-quad", nb.NB_SCENE_UNIFORM_DISC, 1_000_000, nb.NB_TREE_QUAD, ob.KIND_PIPE_QUAD, 200_000),
    ("plummer-4M-lbvh", nb.NB_SCENE_PLUMMER, 4_000_000, nb.NB_TREE_LBVH, ob.KIND_PIPE_LBVH, 4_000),
    ("galaxy-16M-quad", nb.NB_SCENE_TWO_GALAXY, 16_000_000, nb.NB_TREE_QUAD, ob.KIND_PIPE_QUAD, 20_000),
]


@pytest.mark.parametrize("name,scene,n,tree,okind,nsample", CASES, ids=[c[0] for c in CASES])
def test_full_size_parity(name, scene, n, tree, okind, nsample):
    theta = 0.5
    p = nb.generate_scene(scene, 1, n)
    w = nb.World(len(p), theta=theta, tree=tree, force_mode=nb.NB_FORCE_EXACT)
    w.set_particles(p)
    c = nb.Counting()
    ax, ay = w.accel(c)
    g = w.tree()
    T = ob.Tree(okind, p["x"], p["y"], p["weight"])
    o = T.export()
    # the whole tree, bit for bit
    assert g["info"].n_nodes == o["info"].n_nodes and g["info"].overflow_leaves == o["info"].overflow_leaves
    for k in ("keys", "order", "node_depth", "node_start", "node_count", "node_skip", "node_parent", "node_leaf",
              "node_mass", "node_cogx", "node_cogy", "node_x0", "node_y0", "node_x1", "node_y1", "node_s2"):
        assert np.array_equal(bits(g[k]), bits(o[k])), f"{name}: {k}"
    # size-independent invariants
    assert int(g["node_mass"][0]) == int(p["weight"].astype(np.uint64).sum() % (1 << 32))
    order = g["order"]
    seen = np.zeros(len(p), np.uint8)
    seen[order] = 1
    assert seen.all()                                             # a permutation
    nin = int(g["info"].n_in_tree)
    keys_sorted = g["keys"][order[:nin]]
    leaf = g["node_leaf"] == 1
    assert int(g["node_count"][leaf].sum()) == nin               # leaves tile the bodies
    starts = g["node_start"][leaf]
    firstkey = keys_sorted[starts]
    assert (np.diff(firstkey.astype(np.int64)) > 0).all()        # leaves in strictly increasing Morton order
    # forces on a sample: bit-exact against the oracle walk
    sample = np.linspace(0, len(p) - 1, nsample).astype(np.uint32)
    rx, ry, _, _, _ = T.accel(theta, sample)
    assert np.array_equal(bits(ax[sample]), bits(rx)) and np.array_equal(bits(ay[sample]), bits(ry))
    # FAST vs EXACT over all bodies
    w.set_params(force_mode=nb.NB_FORCE_FAST)
    cf = nb.Counting()
    fx, fy = w.accel(cf)
    err = ob.normwise_err(fx, fy, ax, ay)
    assert err <= 1e-5, err
    assert abs(int(cf.interactions) - int(c.interactions)) <= 1e-5 * c.interactions
    w.close()


def test_direct_sum_momentum_and_tree_error_at_1M():
    # sum_i m_i a_i = 0 for the pairwise-antisymmetric force law; tree error vs direct at theta = 0.5
    p = nb.generate_scene(nb.NB_SCENE_UNIFORM_DISC, 2, 1_000_000)
    w = nb.World(len(p), theta=0.5)
    w.set_particles(p)
    ax, ay = w.accel()
    first, cnt = 400_000, 16_384
    dx, dy, _ = w.accel_direct(first, cnt)
    err = ob.normwise_err(ax[first:first + cnt], ay[first:first + cnt], dx, dy)
    assert err < 2e-3, err
    rx, ry = ob.direct_f64(p["x"], p["y"], p["weight"], np.arange(first, first + 512))
    assert ob.normwise_err(dx[:512], dy[:512], rx, ry) <= 2e-6
    w.close()


def test_full_size_reference_bvh_4M_plummer():
    # BASELINE.json configs[2] with the reference's OWN tree (BVHTree::from): whole tree bit for bit, forces on a
    # sample bit for bit against the literal recursive walk.  (At theta = 0.5 this tree opens almost everything on a
    # concentrated scene — ~3M interactions per body — so the force check runs at the reference's shipped theta = 50.)
    p = nb.generate_scene(nb.NB_SCENE_PLUMMER, 1, 4_000_000)
    w = nb.World(len(p), theta=50.0, tree=nb.NB_TREE_BVH_REF, force_mode=nb.NB_FORCE_EXACT)
    w.set_particles(p)
    c = nb.Counting()
    ax, ay = w.accel(c)
    g = w.tree()
    T = ob.Tree(ob.KIND_LITERAL_BVH, p["x"], p["y"], p["weight"])
    o = T.export()
    assert g["info"].n_nodes == o["info"].n_nodes and g["info"].max_depth == o["info"].max_depth
    for k in ("order", "node_depth", "node_start", "node_count", "node_skip", "node_parent", "node_leaf",
              "node_mass", "node_cogx", "node_cogy", "node_x0", "node_y0", "node_x1", "node_y1", "node_s2"):
        assert np.array_equal(bits(g[k]), bits(o[k])), k
    leaf = g["node_leaf"] == 1
    assert int(g["node_count"][leaf].sum()) == len(p) and int(g["node_count"][leaf].max()) <= 64
    sample = np.linspace(0, len(p) - 1, 20_000).astype(np.uint32)
    rx, ry, _, _, _ = T.accel(50.0, sample, literal=True)
    assert np.array_equal(bits(ax[sample]), bits(rx)) and np.array_equal(bits(ay[sample]), bits(ry))
    w.close()


@pytest.mark.parametrize("tree", [nb.NB_TREE_QUAD, nb.NB_TREE_BVH_REF])
def test_mass_linearity_is_exact_at_1M(tree):
    # doubling every weight doubles every u32 node mass, leaves every centre of gravity unchanged and doubles
    # every pair term: scaling by 2 commutes with f32 rounding, so accelerations double BIT FOR BIT (no oracle)
    p = nb.generate_scene(nb.NB_SCENE_UNIFORM_DISC, 5, 1_000_000)
    theta = 0.5 if tree == nb.NB_TREE_QUAD else 50.0
    w = nb.World(len(p), theta=theta, tree=tree, force_mode=nb.NB_FORCE_FAST)
    w.set_particles(p)
    ax, ay = w.accel()
    q = p.copy()
    q["weight"] *= 2
    w.set_particles(q)
    bx, by = w.accel()
    assert np.array_equal(bits(np.float32(2.0) * ax), bits(bx)) and np.array_equal(bits(np.float32(2.0) * ay), bits(by))
    w.close()
