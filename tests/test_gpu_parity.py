"""GPU parity tests: the CUDA path (through the C ABI) against the CPU oracle on the same seeded
inputs.  Bit-exact for Morton keys, sort order, topology, masses, centres of gravity, bounds and —
in NB_FORCE_EXACT mode — forces and trajectories; stated normwise tolerance for NB_FORCE_FAST."""
import numpy as np
import pytest

import nbody_simulation_b200 as nb
import oracle_binding as ob

pytestmark = pytest.mark.gpu

TREE_KEYS = ["keys", "order", "node_depth", "node_start", "node_count", "node_skip", "node_parent", "node_leaf",
             "node_mass", "node_cogx", "node_cogy", "node_x0", "node_y0", "node_x1", "node_y1", "node_s2"]


def bits(a):
    return np.ascontiguousarray(a).view(np.uint8)


def scene(name, n, seed=1):
    kinds = {"default": nb.NB_SCENE_DEFAULT, "disc": nb.NB_SCENE_UNIFORM_DISC, "plummer": nb.NB_SCENE_PLUMMER,
             "galaxy": nb.NB_SCENE_TWO_GALAXY}
    return nb.generate_scene(kinds[name], seed, n)


def assert_tree_equal(g, o, what):
    assert g["info"].n_nodes == o["info"].n_nodes, what
    assert g["info"].n_in_tree == o["info"].n_in_tree, what
    assert g["info"].n_leaves == o["info"].n_leaves, what
    assert g["info"].max_depth == o["info"].max_depth, what
    assert g["info"].overflow_leaves == o["info"].overflow_leaves, what
    for k in TREE_KEYS:
        same = np.array_equal(bits(g[k]), bits(o[k]))
        if not same:
            bad = np.nonzero(g[k] != o[k])[0]
            raise AssertionError(f"{what}: {k} differs at {len(bad)} of {len(g[k])} entries, first {bad[:5]}: "
                                 f"gpu {g[k][bad[:5]]} oracle {o[k][bad[:5]]}")


@pytest.mark.parametrize("tree,okind", [(nb.NB_TREE_QUAD, ob.KIND_PIPE_QUAD), (nb.NB_TREE_LBVH, ob.KIND_PIPE_LBVH)])
@pytest.mark.parametrize("sc,n", [("disc", 100_000), ("default", 0), ("plummer", 150_000), ("galaxy", 120_001)])
def test_tree_bitexact(tree, okind, sc, n):
    p = scene(sc, n)
    w = nb.World(len(p), tree=tree)
    w.set_particles(p)
    w.build_tree()
    g = w.tree()
    o = ob.Tree(okind, p["x"], p["y"], p["weight"]).export()
    assert_tree_equal(g, o, f"{sc}/{tree}")
    if tree == nb.NB_TREE_QUAD:  # and against the literal insert-built quad_tree.rs
        lit = ob.Tree(ob.KIND_LITERAL_QUAD, p["x"], p["y"], p["weight"]).export()
        lit["keys"] = o["keys"]
        assert_tree_equal(g, lit, f"{sc}/literal quad_tree.rs")
    w.close()


@pytest.mark.parametrize("tree,okind,cap", [(nb.NB_TREE_QUAD, ob.KIND_PIPE_QUAD, 1), (nb.NB_TREE_QUAD, ob.KIND_PIPE_QUAD, 3),
                                            (nb.NB_TREE_QUAD, ob.KIND_PIPE_QUAD, 33), (nb.NB_TREE_QUAD, ob.KIND_PIPE_QUAD, 64),
                                            (nb.NB_TREE_LBVH, ob.KIND_PIPE_LBVH, 1), (nb.NB_TREE_LBVH, ob.KIND_PIPE_LBVH, 7),
                                            (nb.NB_TREE_LBVH, ob.KIND_PIPE_LBVH, 40)])
def test_tree_bitexact_other_leaf_caps(tree, okind, cap):
    # leaf_cap is a runtime field (MAX_CAPACITY quad_tree.rs:54 / TARGET_POINTS bvh_tree.rs:37 are consts in the
    # reference): the window of the count kernel, the register / warp / generic leaf fix-up paths all depend on it
    p = scene("galaxy", 70_001, seed=5)
    w = nb.World(len(p), tree=tree, leaf_cap=cap)
    w.set_particles(p)
    w.build_tree()
    assert_tree_equal(w.tree(), ob.Tree(okind, p["x"], p["y"], p["weight"], cap=cap).export(), f"cap {cap}/{tree}")
    w.close()  # (the literal insert-built oracle tree uses the reference's constant capacity: compared at cap 8 above)


@pytest.mark.parametrize("tree,okind", [(nb.NB_TREE_QUAD, ob.KIND_PIPE_QUAD), (nb.NB_TREE_LBVH, ob.KIND_PIPE_LBVH)])
@pytest.mark.parametrize("theta", [0.5, 50.0])
def test_accel_exact_is_bit_identical(tree, okind, theta):
    p = scene("default", 0)
    w = nb.World(len(p), theta=theta, tree=tree, force_mode=nb.NB_FORCE_EXACT)
    w.set_particles(p)
    c = nb.Counting()
    ax, ay = w.accel(c)
    T = ob.Tree(okind, p["x"], p["y"], p["weight"])
    rx, ry, ni, nt, _ = T.accel(theta)
    assert c.interactions == ni and c.node_tests == nt
    assert np.array_equal(bits(ax), bits(rx)) and np.array_equal(bits(ay), bits(ry))
    w.close()


# theta 0.5 and 0.7: Rectangle::contains elided (theta^2 <= 0.49, traverse.cuh kThetaSqNoContains); 0.9: kept
@pytest.mark.parametrize("theta", [0.5, 0.7, 0.9])
@pytest.mark.parametrize("tree,okind", [(nb.NB_TREE_QUAD, ob.KIND_PIPE_QUAD), (nb.NB_TREE_LBVH, ob.KIND_PIPE_LBVH)])
def test_accel_fast_within_tolerance(tree, okind, theta):
    # tolerance: normwise ||da||2/||a||2 <= 1e-5 against the oracle on the SAME tree and theta
    p = scene("disc", 200_000)
    w = nb.World(len(p), theta=theta, tree=tree, force_mode=nb.NB_FORCE_FAST)
    w.set_particles(p)
    c = nb.Counting()
    ax, ay = w.accel(c)
    T = ob.Tree(okind, p["x"], p["y"], p["weight"])
    rx, ry, ni, nt, _ = T.accel(theta)
    # FAST evaluates the opening test as s2 < fma(dy,dy,dx*dx)*theta^2: decisions may flip only when
    # the two sides agree to ~1e-7, so the interaction lists agree to a few parts per million
    assert abs(int(c.interactions) - ni) <= 1e-5 * ni and abs(int(c.node_tests) - nt) <= 1e-5 * nt
    err = ob.normwise_err(ax, ay, rx, ry)
    assert err <= 1e-5, err
    w.close()


def test_quad_literal_reference_forces():
    # GPU exact forces == literal recursive quad_tree.rs walk (carried-over criterion) bit for bit
    p = scene("disc", 60_000)
    w = nb.World(len(p), theta=0.5, tree=nb.NB_TREE_QUAD, force_mode=nb.NB_FORCE_EXACT)
    w.set_particles(p)
    ax, ay = w.accel()
    T = ob.Tree(ob.KIND_LITERAL_QUAD, p["x"], p["y"], p["weight"])
    rx, ry, _, _, _ = T.accel(0.5, literal=True)
    assert np.array_equal(bits(ax), bits(rx)) and np.array_equal(bits(ay), bits(ry))
    w.close()


@pytest.mark.parametrize("tree,okind", [(nb.NB_TREE_QUAD, ob.KIND_PIPE_QUAD), (nb.NB_TREE_LBVH, ob.KIND_PIPE_LBVH)])
def test_trajectory_exact_is_bit_identical(tree, okind):
    p = scene("default", 0)[:40_000].copy()
    theta, dt, steps = 0.7, 0.1, 6
    w = nb.World(len(p), theta=theta, dt=dt, tree=tree, force_mode=nb.NB_FORCE_EXACT)
    w.set_particles(p)
    c = nb.Counting()
    w.step(steps, c)
    ow = ob.OracleWorld(p)
    oc = None
    for _ in range(steps):
        oc = ow.update(okind, dt, theta, counting=oc)
    x, y, vx, vy, m = ow.get()
    q = w.particles
    assert c.interactions == oc.interactions and c.node_tests == oc.node_tests
    for a, b in ((q["x"], x), (q["y"], y), (q["vx"], vx), (q["vy"], vy), (q["weight"], m)):
        assert np.array_equal(bits(a), bits(b))
    w.close()


def test_world_update_host_roundtrip():
    p = scene("disc", 50_000)
    w = nb.World(len(p), theta=0.5, force_mode=nb.NB_FORCE_EXACT)
    host = p.copy()
    w.update_host(host, 0.1)
    ow = ob.OracleWorld(p)
    ow.update(ob.KIND_PIPE_QUAD, 0.1, 0.5)
    x, y, vx, vy, _ = ow.get()
    assert np.array_equal(bits(host["x"]), bits(x)) and np.array_equal(bits(host["vy"]), bits(vy))
    w.close()


def test_direct_sum_against_fp64():
    p = scene("disc", 30_000)
    w = nb.World(len(p), force_mode=nb.NB_FORCE_FAST)
    w.set_particles(p)
    first, count = 1000, 4096
    ax, ay, _ = w.accel_direct(first, count)
    rx, ry = ob.direct_f64(p["x"], p["y"], p["weight"], np.arange(first, first + count))
    err = ob.normwise_err(ax, ay, rx, ry)
    assert err <= 2e-6, err  # fp32 pair terms, fp32 256-tile partials, fp64 across tiles
    w.close()


def test_tree_force_vs_direct_sum():
    # Barnes-Hut error level at theta = 0.5 (not a parity tolerance: documents the BH approximation)
    p = scene("disc", 100_000)
    w = nb.World(len(p), theta=0.5, force_mode=nb.NB_FORCE_FAST)
    w.set_particles(p)
    ax, ay = w.accel()
    dx, dy, _ = w.accel_direct(0, 8192)
    err = ob.normwise_err(ax[:8192], ay[:8192], dx, dy)
    assert err < 5e-3, err
    w.close()


@pytest.mark.parametrize("tree,okind", [(nb.NB_TREE_QUAD, ob.KIND_PIPE_QUAD), (nb.NB_TREE_LBVH, ob.KIND_PIPE_LBVH)])
@pytest.mark.parametrize("n", [1, 2, 8, 9, 64, 65, 1000])
def test_small_and_ragged(tree, okind, n):
    p = scene("disc", n, seed=7)
    w = nb.World(max(n, 1), theta=0.5, tree=tree, force_mode=nb.NB_FORCE_EXACT)
    w.set_particles(p)
    ax, ay = w.accel()
    g = w.tree()
    T = ob.Tree(okind, p["x"], p["y"], p["weight"])
    assert_tree_equal(g, T.export(), f"n={n}")
    rx, ry, _, _, _ = T.accel(0.5)
    assert np.array_equal(bits(ax), bits(rx)) and np.array_equal(bits(ay), bits(ry))
    w.close()


def test_out_of_domain_and_coincident_bodies():
    rng = np.random.default_rng(3)
    n = 5000
    p = np.zeros(n, nb.PARTICLE_DTYPE)
    p["x"] = rng.uniform(-2e4, 1.2e5, n).astype(np.float32)
    p["y"] = rng.uniform(-2e4, 1.2e5, n).astype(np.float32)
    p["weight"] = rng.integers(1, 50, n)
    p["x"][100:140] = 777.25  # 40 coincident bodies: an over-full leaf at the depth cap
    p["y"][100:140] = 4242.5
    p["x"][200] = 0.0  # on the root's lower edge: inside (within_bounds uses >=)
    p["x"][201] = 100000.0  # on the upper edge: outside
    for tree, okind in ((nb.NB_TREE_QUAD, ob.KIND_PIPE_QUAD), (nb.NB_TREE_LBVH, ob.KIND_PIPE_LBVH)):
        w = nb.World(n, theta=0.5, tree=tree, force_mode=nb.NB_FORCE_EXACT)
        w.set_particles(p)
        ax, ay = w.accel()
        g = w.tree()
        T = ob.Tree(okind, p["x"], p["y"], p["weight"])
        o = T.export()
        assert_tree_equal(g, o, f"edge/{tree}")
        if tree == nb.NB_TREE_QUAD:
            assert g["info"].n_in_tree < n and g["info"].overflow_leaves == 1
        rx, ry, _, _, _ = T.accel(0.5)
        assert np.array_equal(bits(ax), bits(rx)) and np.array_equal(bits(ay), bits(ry))
        w.close()


def test_u32_mass_wraps_like_the_reference():
    p = scene("disc", 3000, seed=5)
    p["weight"] = 3_000_000_000  # sums wrap modulo 2^32 exactly like Rust release-mode u32 adds
    w = nb.World(len(p), theta=0.5, force_mode=nb.NB_FORCE_EXACT)
    w.set_particles(p)
    w.build_tree()
    g = w.tree()
    o = ob.Tree(ob.KIND_PIPE_QUAD, p["x"], p["y"], p["weight"]).export()
    assert_tree_equal(g, o, "wrap")
    w.close()


def test_lbvh_auto_domain():
    p = scene("plummer", 50_000)
    w = nb.World(len(p), theta=0.5, tree=nb.NB_TREE_LBVH, domain_auto=True, force_mode=nb.NB_FORCE_EXACT)
    w.set_particles(p)
    ax, ay = w.accel()
    g = w.tree()
    T = ob.Tree(ob.KIND_PIPE_LBVH, p["x"], p["y"], p["weight"], domain_auto=True)
    assert_tree_equal(g, T.export(), "auto-domain")
    rx, ry, _, _, _ = T.accel(0.5)
    assert np.array_equal(bits(ax), bits(rx)) and np.array_equal(bits(ay), bits(ry))
    w.close()


@pytest.mark.parametrize("tree,okind", [(nb.NB_TREE_QUAD, ob.KIND_PIPE_QUAD), (nb.NB_TREE_LBVH, ob.KIND_PIPE_LBVH)])
@pytest.mark.parametrize("domain", [(-20000.0, 12345.678, 123456.7), (1.0e6, 1.0e6, 1.0e5), (16777216.0, 16777216.0, 4096.0)])
def test_other_fixed_domains(tree, okind, domain):
    # root squares other than the reference's (0, 0, 100000): keys and trees must equal the oracle's also where the
    # f32 grid is coarser than the deep cells (offsets of 2^24 with a 4096-wide square) and with bodies outside
    x0, y0, h = domain
    rng = np.random.default_rng(11)
    n = 60_000
    p = np.zeros(n, dtype=scene("disc", 2).dtype)
    p["x"] = (x0 + h * rng.uniform(-0.05, 1.05, n)).astype(np.float32)
    p["y"] = (y0 + h * rng.uniform(-0.05, 1.05, n)).astype(np.float32)
    p["weight"] = rng.integers(1, 5, n)
    w = nb.World(n, tree=tree, domain=domain)
    w.set_particles(p)
    w.build_tree()
    o = ob.Tree(okind, p["x"], p["y"], p["weight"], domain=domain).export()
    assert_tree_equal(w.tree(), o, f"domain {domain}/{tree}")
    w.close()


def test_snapshot_resume_is_bit_identical(tmp_path):
    p = scene("galaxy", 30_000)
    a = nb.World(len(p), theta=0.5, force_mode=nb.NB_FORCE_EXACT)
    a.set_particles(p)
    a.step(3)
    a.save(tmp_path / "s.nbb", 3)
    a.step(2)
    b = nb.World(len(p), theta=0.5, force_mode=nb.NB_FORCE_EXACT)
    assert b.load(tmp_path / "s.nbb") == 3
    b.step(2)
    assert np.array_equal(bits(a.particles), bits(b.particles))
    a.close()
    b.close()


def test_draw_matches_sequential_reference():
    p = scene("default", 0)
    w = nb.World(len(p))
    w.set_particles(p)
    frame = w.draw()
    # sequential restatement of draw() (src/main.rs:41-72)
    H, R = 100000, 1250
    ref = np.zeros((R * R, 4), np.uint8)
    for q in p:
        x, y = q["x"], q["y"]
        if not (y < H and x < H and y >= 0 and x >= 0):
            continue
        off = (int(y) // (H // R)) * R + (int(x) // (H // R))
        if q["weight"] > 10:
            ref[off] = (0, 255, 0, 255)
        elif ref[off, 3] != 255:
            sp = np.float32(np.float32(abs(q["vx"]) + abs(q["vy"])) * np.float32(10.0))
            vel = 0x10 + min(int(min(max(sp, 0), 255)), 0xef)
            ref[off, 0] = 255
            ref[off, 1] = 255 - vel
            ref[off, 2] = 255 - vel
            if ref[off, 3] <= 240:
                ref[off, 3] += 10
    assert np.array_equal(frame.reshape(-1, 4), ref)
    w.close()


@pytest.mark.parametrize("sc,kind", [("disc", nb.NB_SCENE_UNIFORM_DISC), ("plummer", nb.NB_SCENE_PLUMMER)])
def test_device_scene_generation_matches_host(sc, kind):
    # same counter-based draws; the host's and the device's fp64 sin/cos may differ across an f32 rounding
    # boundary for ~1 value in 1e9: allow one ulp on at most 5 of 2M coordinates, everything else bit-equal
    n = 1_000_000
    p = scene(sc, n, seed=7)
    w = nb.World(n)
    w.generate_on_device(kind, 7, n)
    q = w.particles
    for k in ("x", "y"):
        d = np.abs(q[k].view(np.int32).astype(np.int64) - p[k].view(np.int32).astype(np.int64))
        assert d.max() <= 1 and int((d != 0).sum()) <= 5, (k, int(d.max()), int((d != 0).sum()))
    assert np.array_equal(q["weight"], p["weight"]) and not q["vx"].any() and not q["vy"].any()
    w.close()


@pytest.mark.parametrize("sc,kind,n", [("galaxy", nb.NB_SCENE_TWO_GALAXY, 300_001), ("galaxy", nb.NB_SCENE_TWO_GALAXY, 2),
                                       ("default", nb.NB_SCENE_DEFAULT, 0)])
def test_device_generation_of_the_sequentially_drawn_scenes(sc, kind, n):
    # two-galaxy (rejection sampling) and World::new's scene (grid acceptance): the same bodies in the same order as
    # the host's sequential loops; fp64 sin/cos (galaxy) and the default scene's f32 cosf/sinf may differ by one ulp
    p = scene(sc, n, seed=5)
    w = nb.World(max(len(p), 200_000))
    w.generate_on_device(kind, 5, n)
    q = w.particles
    assert len(q) == len(p)
    assert np.array_equal(q["weight"], p["weight"])
    for k in ("x", "y", "vx", "vy"):
        d = np.abs(q[k].view(np.int32).astype(np.int64) - p[k].view(np.int32).astype(np.int64))
        # glibc's f32 cosf / sinf (default scene, rand_disc) are not correctly rounded: ~1 % of those values sit 1-2 ulp
        # from the device's fp64-then-round results; the fp64 draws (galaxy) differ for a handful at most
        lim = 2e-2 * len(p) if sc == "default" else 1e-4 * len(p) + 5
        assert d.max() <= 2 and int((d != 0).sum()) <= lim, (k, int(d.max()), int((d != 0).sum()))
    w.close()


# ---- top-digit (speculative) sort of the quadtree build: radix_sort.cuh ---------------------------------------
@pytest.mark.parametrize("sc,n", [("disc", 200_000), ("plummer", 150_000), ("galaxy", 120_001), ("default", 0)])
def test_top_digit_sort_builds_the_same_tree(sc, n):
    # the first build of a context sorts every digit (no depth known yet); the second sorts only the top digits the
    # first one's depth calls for and checks the guess on the device: same tree, bit for bit, as the oracle's
    p = scene(sc, n)
    o = ob.Tree(ob.KIND_PIPE_QUAD, p["x"], p["y"], p["weight"]).export()
    w = nb.World(len(p))
    w.set_particles(p)
    w.build_tree()
    first = w.tree_info()
    assert first.sort_top_digits == 0
    w.set_particles(p)
    w.build_tree()
    info = w.tree_info()
    # one level of margin, two key bits per level below the outside bit, digits of 8 or 9 bits (make_top_plan)
    bits_needed = 2 * max(first.max_depth + 1, 8) + 1
    want = -(-bits_needed // 9)
    cover = max(bits_needed, 8 * want)
    if want >= 6 or cover >= 49:
        want, cover = 0, 1
    assert info.sort_top_digits == want and info.sort_top_levels == (cover - 1) // 2 and info.sort_redone == 0, \
        (want, cover, info.sort_top_digits, info.sort_top_levels)
    assert_tree_equal(w.tree(), o, f"{sc}/top-digit sort")
    f = nb.World(len(p), full_sort=True)
    f.set_particles(p)
    f.build_tree()
    f.set_particles(p)
    f.build_tree()
    assert f.tree_info().sort_top_digits == 0
    assert_tree_equal(f.tree(), o, f"{sc}/full sort")
    w.close()
    f.close()


@pytest.mark.parametrize("levels,widths", [(13, 3), (14, 4), (17, 4), (18, 5), (22, 5)])
def test_top_digit_sort_every_digit_width(monkeypatch, levels, widths):
    # plans with several 9-bit digits only appear at particular depths (27 / 35 / 45 key bits): ask for that many levels
    monkeypatch.setenv("NB_B200_SORT_MIN_LEVELS", str(levels))
    p = scene("galaxy", 150_001)
    o = ob.Tree(ob.KIND_PIPE_QUAD, p["x"], p["y"], p["weight"]).export()
    w = nb.World(len(p))
    w.set_particles(p)
    w.build_tree()
    w.set_particles(p)
    w.build_tree()
    info = w.tree_info()
    assert info.sort_redone == 0 and info.sort_top_digits == widths and info.sort_top_levels >= levels, \
        (info.sort_top_digits, info.sort_top_levels)
    assert_tree_equal(w.tree(), o, f"top-digit sort over {levels} levels")
    w.close()


# ---- repairing the previous order instead of sorting: sort_repair.cuh ------------------------------------------
def _moving(name, n, movers_per_step, seed=5):
    """A scene whose bodies drift in random directions, fast enough for roughly `movers_per_step` of them to leave the
    cell of their leaf in a step of dt = 0.1 (0: at rest) — fewer in fact: the estimate takes every leaf to be as small
    as the deepest ones."""
    p = scene(name, n)
    w = nb.World(len(p))
    w.set_particles(p)
    w.build_tree()
    info = w.tree_info()
    w.close()
    cell = info.root_h / 2.0 ** info.max_depth
    speed = movers_per_step / (1.27 * len(p)) * cell / 0.1  # a body moving d crosses a cell line with p = 1.27 d / cell
    rng = np.random.default_rng(seed)
    ang = rng.uniform(0, 2 * np.pi, len(p))
    p["vx"] = (speed * np.cos(ang)).astype(np.float32)
    p["vy"] = (speed * np.sin(ang)).astype(np.float32)
    return p


@pytest.mark.parametrize("sc,n,movers", [("disc", 200_000, 0), ("galaxy", 150_001, 900), ("disc", 300_000, 1500),
                                         ("disc", 5_000, 300)])
def test_sort_repair_builds_the_same_tree(monkeypatch, sc, n, movers):
    # a build whose state is still in the previous tree's order counts the bodies that left their cell; once that count
    # is small the builds stop sorting and only re-insert those bodies.  Tree and trajectory stay bit-identical to the
    # oracle's / to a context that always sorts
    p = _moving(sc, n, movers)
    w = nb.World(len(p), theta=0.5, dt=0.1, force_mode=nb.NB_FORCE_EXACT)
    w.set_particles(p)
    w.step(1)          # full sort (no depth known)
    w.step(2)          # top-digit sorts that count the movers
    w.step(2)          # repairs
    w.build_tree()
    info = w.tree_info()
    assert info.sort_redone == 0 and info.sort_top_digits > 0
    assert info.sort_repaired == 1, ("the build after a step must repair the order", info.sort_movers, info.n_in_tree, info.n_bodies, info.overflow_leaves)
    if movers == 0:
        assert info.sort_movers <= 20, info.sort_movers
    else:
        assert 0 < info.sort_movers <= 2048, info.sort_movers
    q = w.particles
    o = ob.Tree(ob.KIND_PIPE_QUAD, q["x"], q["y"], q["weight"]).export()
    assert_tree_equal(w.tree(), o, f"{sc}/repaired order, {info.sort_movers} movers")
    monkeypatch.setenv("NB_B200_NO_REPAIR", "1")
    ref = nb.World(len(p), theta=0.5, dt=0.1, force_mode=nb.NB_FORCE_EXACT)
    ref.set_particles(p)
    ref.step(1)
    ref.step(4)
    ref.build_tree()
    assert ref.tree_info().sort_repaired == 0
    assert np.array_equal(bits(q), bits(ref.particles))
    c, cr = nb.Counting(), nb.Counting()
    w.step(5, c)
    ref.step(5, cr)
    assert c.interactions == cr.interactions and c.node_tests == cr.node_tests
    assert np.array_equal(bits(w.particles), bits(ref.particles))
    w.close()
    ref.close()


def test_sort_repair_gives_up_when_too_many_bodies_moved(monkeypatch):
    # the movers were few when counted, then a burst of them: more than a repair takes.  The step is redone with the
    # radix sort (inside a batch of steps) and the trajectory is that of a context that always sorts; a scene that
    # moves a lot from the start never tries
    p = _moving("disc", 100_000, 600)
    w = nb.World(len(p), theta=0.5, dt=0.1, force_mode=nb.NB_FORCE_EXACT)
    w.set_particles(p)
    w.step(1)
    w.step(2)
    fast = w.particles
    fast["vx"] *= 200.0
    fast["vy"] *= 200.0
    ref_start = fast.copy()
    w.set_particles(fast)  # an upload: in caller order, so no repair until the order is the tree's again
    w.step(1)
    w.step(1)
    w.build_tree()
    info = w.tree_info()
    assert info.sort_repaired == 0 and info.sort_redone == 0  # counted ~40 000 movers: stays with the radix sort
    w.close()
    # the burst itself: slow bodies, repairs running, then the velocities jump on the device side of a batch?  There
    # is no such call — so shrink the cells instead: a context whose repair limit is hit because dt grows
    w = nb.World(len(p), theta=0.5, dt=0.1, force_mode=nb.NB_FORCE_EXACT)
    w.set_particles(p)
    w.step(1)
    w.step(2)          # counted ~200 movers
    w.set_params(0.5, 40.0, nb.NB_FORCE_EXACT)  # dt x400: ~80 000 movers in the next step
    c = nb.Counting()
    w.step(4, c)       # the first repair fails; the batch is redone with the radix sort
    assert c.steps == 4
    monkeypatch.setenv("NB_B200_NO_REPAIR", "1")
    ref = nb.World(len(p), theta=0.5, dt=0.1, force_mode=nb.NB_FORCE_EXACT)
    ref.set_particles(p)
    ref.step(1)
    ref.step(2)
    ref.set_params(0.5, 40.0, nb.NB_FORCE_EXACT)
    cr = nb.Counting()
    ref.step(4, cr)
    assert c.interactions == cr.interactions
    assert np.array_equal(bits(w.particles), bits(ref.particles))
    w.close()
    ref.close()


def test_top_digit_sort_falls_back_when_the_guess_is_too_shallow():
    shallow, deep = scene("disc", 20_000), scene("plummer", 150_000)
    w = nb.World(len(deep), theta=0.5, force_mode=nb.NB_FORCE_EXACT)
    w.set_particles(shallow)
    w.build_tree()
    d0 = w.tree_info().max_depth
    w.set_particles(deep)  # much deeper than the previous build: the guess (from d0) cannot hold
    c = nb.Counting()
    ax, ay = w.accel(c)
    info = w.tree_info()
    assert info.max_depth > info.sort_top_levels >= d0, (d0, info.sort_top_levels, info.max_depth)
    assert info.sort_redone == 1
    T = ob.Tree(ob.KIND_PIPE_QUAD, deep["x"], deep["y"], deep["weight"])
    assert_tree_equal(w.tree(), T.export(), "fallback to the full sort")
    rx, ry, ni, nt, _ = T.accel(0.5)
    assert int(c.interactions) == ni and int(c.node_tests) == nt
    assert np.array_equal(bits(ax), bits(rx)) and np.array_equal(bits(ay), bits(ry))
    w.set_particles(deep)  # now the depth is known: the guess holds
    w.build_tree()
    info = w.tree_info()
    assert info.sort_redone == 0 and info.sort_top_levels >= info.max_depth
    assert_tree_equal(w.tree(), T.export(), "after the fallback")
    w.close()


# ---- round 2 ---------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("theta", [0.5, 0.9])
@pytest.mark.parametrize("tree,okind", [(nb.NB_TREE_QUAD, ob.KIND_PIPE_QUAD), (nb.NB_TREE_LBVH, ob.KIND_PIPE_LBVH)])
def test_group_walk_makes_the_same_decisions(tree, okind, theta):
    # nb_config.fast_walk_group: lane-parallel decisions + staged lists; every body must still take exactly the
    # interactions of its private walk (counts equal the depth-first FAST kernel's), sums within 1e-5 of the oracle
    p = scene("galaxy", 150_001)  # ragged last warp
    res = []
    for grp in (0, 1, 2):  # depth-first, group walk, hybrid (group far field + depth-first near field)
        w = nb.World(len(p), theta=theta, tree=tree, force_mode=nb.NB_FORCE_FAST, fast_walk_group=grp)
        w.set_particles(p)
        c = nb.Counting()
        ax, ay = w.accel(c)
        res.append((ax, ay, int(c.interactions), int(c.node_tests)))
        w.close()
    assert res[0][2] == res[1][2] == res[2][2] and res[0][3] == res[1][3] == res[2][3]
    rx, ry, _, _, _ = ob.Tree(okind, p["x"], p["y"], p["weight"]).accel(theta)
    assert ob.normwise_err(res[1][0], res[1][1], rx, ry) <= 1e-5
    assert ob.normwise_err(res[2][0], res[2][1], rx, ry) <= 1e-5


def nonnormal_scene():
    rng = np.random.default_rng(11)
    n = 3000
    p = np.zeros(n, nb.PARTICLE_DTYPE)
    p["x"] = rng.uniform(1e3, 9.9e4, n).astype(np.float32)
    p["y"] = rng.uniform(1e3, 9.9e4, n).astype(np.float32)
    p["weight"] = rng.integers(1, 9, n)
    p["x"][10] = np.inf               # a body at +inf
    p["y"][20] = np.nan               # a body at NaN
    p["x"][30], p["y"][30] = 5.0e4, 5.0e4
    p["x"][31], p["y"][31] = 5.0e4, 5.0e4  # coincident with 30: the self-pair rule (sum == 0 is not normal)
    p["x"][40:52] = 1234.5            # 12 coincident bodies (> MAX_CAPACITY)
    p["y"][40:52] = 4321.0
    p["x"][60], p["y"][60] = 1e-40, 1e-40   # subnormal coordinates, 1e-40 apart from the next
    p["x"][61], p["y"][61] = 2e-40, 1e-40
    return p


@pytest.mark.parametrize("tree,okind", [(nb.NB_TREE_QUAD, ob.KIND_PIPE_QUAD), (nb.NB_TREE_LBVH, ob.KIND_PIPE_LBVH)])
def test_non_normal_inputs_exact(tree, okind):
    # main.rs:238-243: `if !sum.is_normal() return` — zero, subnormal, infinite and NaN separations contribute nothing
    p = nonnormal_scene()
    w = nb.World(len(p), theta=0.5, tree=tree, force_mode=nb.NB_FORCE_EXACT)
    w.set_particles(p)
    c = nb.Counting()
    ax, ay = w.accel(c)
    T = ob.Tree(okind, p["x"], p["y"], p["weight"])
    rx, ry, ni, nt, _ = T.accel(0.5)
    assert c.interactions == ni and c.node_tests == nt
    for g, r in ((ax, rx), (ay, ry)):
        nan_g, nan_r = np.isnan(g), np.isnan(r)
        assert np.array_equal(nan_g, nan_r)  # NaN payloads differ between x86 and sm_100: compare NaN-ness
        assert np.array_equal(bits(g[~nan_g]), bits(r[~nan_r]))
    w.close()


def test_non_normal_inputs_fast_documented_deviation():
    # NB_FORCE_FAST has no is_normal test (include/nbody_b200.h): zero and subnormal separations still contribute
    # nothing measurable (1e-20 seed), but an infinite or NaN SOURCE poisons the sums that meet it.  Bodies outside
    # the root square are not sources in the quadtree, so there the result stays within tolerance of the oracle.
    p = nonnormal_scene()
    w = nb.World(len(p), theta=0.5, tree=nb.NB_TREE_QUAD, force_mode=nb.NB_FORCE_FAST)
    w.set_particles(p)
    ax, ay = w.accel()
    rx, ry, _, _, _ = ob.Tree(ob.KIND_PIPE_QUAD, p["x"], p["y"], p["weight"]).accel(0.5)
    ok = np.isfinite(p["x"]) & np.isfinite(p["y"])
    assert np.isfinite(ax[ok]).all() and np.isfinite(ay[ok]).all()
    assert ob.normwise_err(ax[ok], ay[ok], rx[ok], ry[ok]) <= 1e-5
    w.close()


def test_quad_empty_and_prune_report_what_the_reference_would_return():
    # quad_tree.rs:66-137 on a tree holding this step's bodies; zero-weight regions make Roots of total_mass 0
    p = scene("disc", 60_000)
    p["weight"][(p["x"] < 30000) & (p["y"] < 40000)] = 0   # a massless corner: whole subtrees of mass 0
    p["weight"][:50] = 0
    for pp in (p, scene("galaxy", 20_000)):
        w = nb.World(len(pp), theta=0.5, tree=nb.NB_TREE_QUAD)
        w.set_particles(pp)
        w.build_tree()
        T = ob.Tree(ob.KIND_LITERAL_QUAD, pp["x"], pp["y"], pp["weight"])
        assert T.export()["info"].n_nodes == w.tree_info().n_nodes
        e, pr = T.fresh_stats()
        assert (w.empty(), w.prune()) == (e, pr)
        w.close()


def test_top_digit_sort_too_shallow_inside_a_batch_of_steps():
    # nb_step enqueues its steps blind; a depth guess that fails in step k turns steps k.. into no-ops that leave the
    # state untouched, and the host redoes them (full sort first): same trajectory as a context that always sorts fully
    shallow, deep = scene("disc", 20_000), scene("plummer", 150_000)
    ref = nb.World(len(deep), theta=0.5, force_mode=nb.NB_FORCE_EXACT, full_sort=True)
    ref.set_particles(deep)
    ref.step(4)
    w = nb.World(len(deep), theta=0.5, force_mode=nb.NB_FORCE_EXACT)
    w.set_particles(shallow)
    w.step(2)             # the depth of the shallow scene is now the guess
    w.set_particles(deep)
    c = nb.Counting()
    w.step(4, c)
    assert c.steps == 4
    assert np.array_equal(bits(w.particles), bits(ref.particles))
    w.close()
    ref.close()


def test_upload_then_device_generation_does_not_reuse_the_stale_upload(monkeypatch):
    # an AoS upload leaves its records in the staging buffer for the one-pass re-order of the next build (from 2M bodies;
    # forced on for every size here): any other way of setting the state must drop that claim
    monkeypatch.setenv("NB_B200_AOS_PERMUTE_MIN", "0")
    n = 50_000
    w = nb.World(n, theta=0.5, force_mode=nb.NB_FORCE_EXACT)
    w.set_particles(scene("galaxy", n))           # upload: the staging buffer describes this state ...
    w.generate_on_device(nb.NB_SCENE_UNIFORM_DISC, 1, n)  # ... and no longer this one
    w.step(2)
    ref = nb.World(n, theta=0.5, force_mode=nb.NB_FORCE_EXACT)
    ref.generate_on_device(nb.NB_SCENE_UNIFORM_DISC, 1, n)
    ref.step(2)
    assert np.array_equal(bits(w.particles), bits(ref.particles))
    # and the re-order itself: upload + steps equals the resident path
    p = scene("galaxy", n)
    w.set_particles(p)
    w.step(2)
    monkeypatch.setenv("NB_B200_AOS_PERMUTE_MIN", "4000000000")
    r2 = nb.World(n, theta=0.5, force_mode=nb.NB_FORCE_EXACT)
    r2.set_particles(p)
    r2.step(2)
    assert np.array_equal(bits(w.particles), bits(r2.particles))
    for x in (w, ref, r2):
        x.close()
