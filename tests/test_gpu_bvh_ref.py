"""GPU parity for NB_TREE_BVH_REF: the device build of the reference's LIVE tree, BVHTree::from
(src/bvh_tree.rs:57-96), against the oracle's literal restatement (oracle kind 3) — topology, body
order, masses, centres of gravity and bounds bit-exact; exact-mode forces and multi-step trajectories
bit-exact (the reference's Vec is permuted by every build, main.rs:400, so later steps depend on the
exact order the partitions leave)."""
import numpy as np
import pytest

import nbody_simulation_b200 as nb
import oracle_binding as ob
from test_gpu_parity import TREE_KEYS, bits, scene

pytestmark = pytest.mark.gpu


def assert_bvh_equal(g, o, what):
    for f in ("n_nodes", "n_in_tree", "n_leaves", "max_depth", "overflow_leaves"):
        assert getattr(g["info"], f) == getattr(o["info"], f), (what, f, getattr(g["info"], f), getattr(o["info"], f))
    for k in TREE_KEYS:
        if k == "keys":
            continue  # this tree has no Morton keys
        a, b = g[k], o[k]
        same = np.array_equal(bits(a), bits(b))
        if not same and a.dtype.kind == "f":  # an empty leaf's 0/0: NaN payloads differ between x86 and sm_100
            same = np.array_equal(np.isnan(a), np.isnan(b)) and np.array_equal(bits(a[~np.isnan(a)]), bits(b[~np.isnan(b)]))
        if not same:
            bad = np.nonzero(a != b)[0]
            raise AssertionError(f"{what}: {k} differs at {len(bad)} of {len(a)} entries, first {bad[:5]}: "
                                 f"gpu {a[bad[:5]]} oracle {b[bad[:5]]}")


@pytest.mark.parametrize("sc,n", [("disc", 100_000), ("default", 0), ("plummer", 150_000), ("galaxy", 120_001),
                                  ("disc", 1_000_000)])
def test_bvh_ref_tree_bitexact(sc, n):
    p = scene(sc, n)
    w = nb.World(len(p), tree=nb.NB_TREE_BVH_REF)
    w.set_particles(p)
    w.build_tree()
    g = w.tree()
    o = ob.Tree(ob.KIND_LITERAL_BVH, p["x"], p["y"], p["weight"]).export()
    assert_bvh_equal(g, o, f"{sc}/{n}")
    w.close()


@pytest.mark.parametrize("n", [1, 2, 64, 65, 129, 5000])
def test_bvh_ref_small_and_edge_sizes(n):
    # the reference always splits the root (bvh_tree.rs:90), even for <= 64 bodies; n = 1 leaves one side empty
    p = scene("disc", max(n, 16))[:n].copy()
    w = nb.World(len(p), tree=nb.NB_TREE_BVH_REF)
    w.set_particles(p)
    w.build_tree()
    assert_bvh_equal(w.tree(), ob.Tree(ob.KIND_LITERAL_BVH, p["x"], p["y"], p["weight"]).export(), f"n={n}")
    w.close()


def test_bvh_ref_positions_off_domain_and_negative():
    # the BVH has no domain: negative and far-away coordinates (max seeded 0, bvh_tree.rs:59; mixed-sign sums)
    r = np.random.default_rng(3)
    p = scene("disc", 50_000)
    p["x"] = r.normal(0.0, 3e4, len(p)).astype(np.float32)
    p["y"] = (r.normal(-2e5, 1e4, len(p))).astype(np.float32)
    w = nb.World(len(p), tree=nb.NB_TREE_BVH_REF)
    w.set_particles(p)
    w.build_tree()
    assert_bvh_equal(w.tree(), ob.Tree(ob.KIND_LITERAL_BVH, p["x"], p["y"], p["weight"]).export(), "mixed sign")
    w.close()


@pytest.mark.parametrize("theta", [0.5, 50.0])
def test_bvh_ref_accel_exact_is_bit_identical(theta):
    p = scene("default", 0)
    w = nb.World(len(p), theta=theta, tree=nb.NB_TREE_BVH_REF, force_mode=nb.NB_FORCE_EXACT)
    w.set_particles(p)
    c = nb.Counting()
    ax, ay = w.accel(c)
    T = ob.Tree(ob.KIND_LITERAL_BVH, p["x"], p["y"], p["weight"])
    rx, ry, ni, nt, _ = T.accel(theta, literal=True)  # the recursive main.rs:349-386 on the pointer tree
    assert c.interactions == ni and c.node_tests == nt
    assert np.array_equal(bits(ax), bits(rx)) and np.array_equal(bits(ay), bits(ry))
    w.close()


def test_bvh_ref_accel_fast_within_tolerance():
    # tolerance: normwise ||da||2/||a||2 <= 1e-5 against the literal reference tree at the same theta
    p = scene("disc", 200_000)
    w = nb.World(len(p), theta=0.5, tree=nb.NB_TREE_BVH_REF, force_mode=nb.NB_FORCE_FAST)
    w.set_particles(p)
    ax, ay = w.accel()
    rx, ry, _, _, _ = ob.Tree(ob.KIND_LITERAL_BVH, p["x"], p["y"], p["weight"]).accel(0.5, literal=True)
    err = ob.normwise_err(ax, ay, rx, ry)
    assert err <= 1e-5, err
    w.close()


def test_bvh_ref_trajectory_is_bit_identical_to_world_update():
    # World::update x 6 (main.rs:388-425, identity semantics): the Vec order left by each build feeds the next
    p = scene("default", 0)[:60_000].copy()
    theta, dt, steps = 0.7, 0.1, 6
    w = nb.World(len(p), theta=theta, dt=dt, tree=nb.NB_TREE_BVH_REF, force_mode=nb.NB_FORCE_EXACT)
    w.set_particles(p)
    c = nb.Counting()
    w.step(steps, c)
    ow = ob.OracleWorld(p)
    oc = None
    for _ in range(steps):
        oc = ow.update(ob.KIND_LITERAL_BVH, dt, theta, counting=oc)
    x, y, vx, vy, m = ow.get()
    q = w.particles
    assert c.interactions == oc.interactions and c.node_tests == oc.node_tests
    for a, b in ((q["x"], x), (q["y"], y), (q["vx"], vx), (q["vy"], vy), (q["weight"], m)):
        assert np.array_equal(bits(a), bits(b))
    w.close()


def test_bvh_ref_coincident_bodies_terminate():
    # > 64 coincident bodies: BVHTree::from never separates them (the reference overflows its stack; the oracle
    # stops at depth 4000, the device build at level 96).  Both end in one over-full leaf reached through a chain
    # of roots whose empty sibling leaves poison their centres of gravity (0/0), so nothing on the chain is ever
    # accepted and the forces are the plain sums over the same body order: bit-identical.
    p = scene("disc", 20_000)
    p["x"][5000:5200] = np.float32(41234.5)
    p["y"][5000:5200] = np.float32(56789.25)
    w = nb.World(len(p), theta=0.5, tree=nb.NB_TREE_BVH_REF, force_mode=nb.NB_FORCE_EXACT)
    w.set_particles(p)
    ax, ay = w.accel()
    info = w.tree()["info"]
    assert info.overflow_leaves >= 1
    rx, ry, _, _, _ = ob.Tree(ob.KIND_LITERAL_BVH, p["x"], p["y"], p["weight"]).accel(0.5, literal=True)
    assert np.array_equal(bits(ax), bits(rx)) and np.array_equal(bits(ay), bits(ry))
    w.close()


def test_bvh_ref_resize_and_reuse_context():
    # the blind level count is a hint from the previous build: a context fed a much smaller / larger scene must adapt
    w = nb.World(300_000, tree=nb.NB_TREE_BVH_REF)
    for n in (300_000, 900, 120_000):
        p = scene("plummer", n)
        w.set_particles(p)
        w.build_tree()
        assert_bvh_equal(w.tree(), ob.Tree(ob.KIND_LITERAL_BVH, p["x"], p["y"], p["weight"]).export(), f"reuse n={n}")
    w.close()


def test_bvh_ref_world_update_host_keeps_the_reference_vec_order():
    # nb_world_update with the Vec on the host: like the reference, the Vec comes back permuted by BVHTree::from,
    # so calling it repeatedly follows the same trajectory as the resident nb_step loop and as the oracle's World
    p = scene("default", 0)[:50_000].copy()
    theta, dt, steps = 0.7, 0.1, 4
    w = nb.World(len(p), theta=theta, dt=dt, tree=nb.NB_TREE_BVH_REF, force_mode=nb.NB_FORCE_EXACT)
    host = p.copy()
    for _ in range(steps):
        w.update_host(host, dt)
    ow = ob.OracleWorld(p)
    for _ in range(steps):
        ow.update(ob.KIND_LITERAL_BVH, dt, theta)
    x, y, vx, vy, m = ow.get()  # by original id; the host Vec is in the reference's permuted order: compare as sets

    def canon(a, b, c, d):
        rows = np.stack([a.view(np.uint32), b.view(np.uint32), c.view(np.uint32), d.view(np.uint32)], axis=1)
        return rows[np.lexsort(rows.T[::-1])]

    assert np.array_equal(canon(host["x"], host["y"], host["vx"], host["vy"]), canon(x, y, vx, vy))
    assert not np.array_equal(bits(host["x"]), bits(x))  # and it really is permuted
    w.close()


def test_bvh_ref_faithful_index_bug_matches_the_reference_as_shipped():
    # main.rs:398-423 verbatim: acceleration[idx] (computed in pre-build Vec order) lands on the body that
    # BVHTree::from moved to idx.  Behind cfg.faithful_index_bug the device step reproduces exactly that.
    p = scene("default", 0)[:40_000].copy()
    theta, dt, steps = 50.0, 0.1, 5
    w = nb.World(len(p), theta=theta, dt=dt, tree=nb.NB_TREE_BVH_REF, force_mode=nb.NB_FORCE_EXACT,
                 faithful_index_bug=True)
    w.set_particles(p)
    w.step(steps)
    ow = ob.OracleWorld(p)
    for _ in range(steps):
        ow.update(ob.KIND_LITERAL_BVH, dt, theta, faithful_bug=True)
    x, y, vx, vy, _ = ow.get()
    q = w.particles
    for a, b in ((q["x"], x), (q["y"], y), (q["vx"], vx), (q["vy"], vy)):
        assert np.array_equal(bits(a), bits(b))
    # and it is observably different from the identity-order step
    ref = nb.World(len(p), theta=theta, dt=dt, tree=nb.NB_TREE_BVH_REF, force_mode=nb.NB_FORCE_EXACT)
    ref.set_particles(p)
    ref.step(steps)
    assert not np.array_equal(bits(ref.particles["vx"]), bits(q["vx"]))
    w.close()
    ref.close()


def test_bvh_ref_snapshot_resume_is_bit_identical(tmp_path):
    # BVHTree::from depends on the Vec order the previous build left: the snapshot keeps slot order + ids (version 2)
    p = scene("galaxy", 30_000)
    a = nb.World(len(p), theta=0.7, tree=nb.NB_TREE_BVH_REF, force_mode=nb.NB_FORCE_EXACT)
    a.set_particles(p)
    a.step(3)
    a.save(tmp_path / "s.nbb", 3)
    a.step(2)
    b = nb.World(len(p), theta=0.7, tree=nb.NB_TREE_BVH_REF, force_mode=nb.NB_FORCE_EXACT)
    assert b.load(tmp_path / "s.nbb") == 3
    b.step(2)
    assert np.array_equal(bits(a.particles), bits(b.particles))  # by original id
    ha, hb = a.particles.copy(), b.particles.copy()
    a.close()
    b.close()
    assert np.array_equal(bits(ha), bits(hb))


def draw_sequential(p, H=100000, R=1250):
    """Sequential restatement of draw() (src/main.rs:41-72) over the Vec order given."""
    ref = np.zeros((R * R, 4), np.uint8)
    for q in p:
        x, y = q["x"], q["y"]
        if not (y < H and x < H and y >= 0 and x >= 0):
            continue
        off = (int(y) // (H // R)) * R + (int(x) // (H // R))
        if q["weight"] > 10:
            ref[off] = (0, 0xFF, 0, 0xFF)
        elif ref[off, 3] != 0xFF:
            sp = (abs(np.float32(q["vx"])) + abs(np.float32(q["vy"]))) * np.float32(10.0)
            v = 0 if np.isnan(sp) else int(min(max(sp, 0), 255))
            vel = 0x10 + min(v, 0xEF)
            a = int(ref[off, 3])
            ref[off] = (0xFF, 0xFF - vel, 0xFF - vel, a + 10 if a <= 240 else a)
    return ref


def test_bvh_ref_draw_follows_the_permuted_vec_order():
    # after BVHTree::from the reference draws its Vec in the post-partition order (main.rs:46 after :400): pixels hit
    # by several light bodies take the colour of the last one IN THAT ORDER
    p = scene("disc", 120_000)
    p["vx"] = (np.arange(len(p)) % 23).astype(np.float32) * 0.5  # distinct colours per body
    w = nb.World(len(p), theta=50.0, dt=0.1, tree=nb.NB_TREE_BVH_REF, force_mode=nb.NB_FORCE_EXACT)
    host = p.copy()
    w.update_host(host, 0.0)  # delta 0: positions and velocities stay, the Vec comes back permuted
    assert not np.array_equal(bits(host["x"]), bits(p["x"]))
    frame = w.draw().reshape(-1, 4)
    assert np.array_equal(frame, draw_sequential(host))
    w.close()


def test_bvh_ref_blind_level_count_is_checked_on_the_device_and_redone():
    # round 2: after the first build the host enqueues the level-synchronous part blind (previous level count + 1, no
    # synchronisation); a deeper tree than guessed turns the step into a no-op and the host redoes it.  Same trajectory
    # as a context that never guessed, including in the middle of an nb_step batch.
    small, big = scene("disc", 20_000), scene("plummer", 120_000)
    ref = nb.World(len(big), theta=0.7, tree=nb.NB_TREE_BVH_REF, force_mode=nb.NB_FORCE_EXACT)
    ref.set_particles(big)
    ref.step(3)
    w = nb.World(len(big), theta=0.7, tree=nb.NB_TREE_BVH_REF, force_mode=nb.NB_FORCE_EXACT)
    w.set_particles(small)
    w.step(2)  # few big levels: that count is the next guess
    w.set_particles(big)
    c = nb.Counting()
    w.step(3, c)
    assert c.steps == 3
    assert np.array_equal(bits(w.particles), bits(ref.particles))
    # and back to a shallow scene: the guess overshoots (no-op levels), nothing to redo
    w.set_particles(small)
    w.build_tree()
    assert_bvh_equal(w.tree(), ob.Tree(ob.KIND_LITERAL_BVH, small["x"], small["y"], small["weight"]).export(), "after overshoot")
    w.close()
    ref.close()
