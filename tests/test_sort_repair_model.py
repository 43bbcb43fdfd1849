"""The counting argument of csrc/sort_repair.cuh, restated in numpy and checked on the CPU.

The CUDA kernels themselves are tested on a GPU (tests/test_gpu_parity.py::test_sort_repair_*: trees and trajectories
bit-identical to the radix-sort path).  This file checks the ALGORITHM they implement, rule by rule as the kernels apply
it — keepers rank inside their old leaf, movers are sorted and merged by counting, each tile of slots places the movers
whose key falls into its range — on real quadtrees (the oracle's) and motions that stress it: movers into gaps, before
the first and behind the last body, into full leaves, across tile boundaries, bursts, nobody moving.  The result must be
a permutation of the slots whose keys are sorted on the prefix the top-digit sort would cover (radix_sort.cuh)."""
import bisect

import numpy as np
import pytest

import oracle_binding as ob

KEY_BITS = 48
H = 32  # kRepairLeafScan


def rp_shift(L):
    return KEY_BITS - 2 * int(min(L, 24))


def old_tree(x, y, w):
    """What the device still holds at the next build: sorted keys and the leaf depth of every sorted position; and the
    body in every slot (the state is in tree order: by id inside a leaf)."""
    t = ob.Tree(ob.KIND_PIPE_QUAD, x, y, w).export()
    n = len(x)
    prev = np.sort(t["keys"])
    ldepth = np.zeros(n, np.int64)
    for i in np.nonzero(t["node_leaf"])[0]:
        s, c = int(t["node_start"][i]), int(t["node_count"][i])
        ldepth[s:s + c] = int(t["node_depth"][i])
    return prev, ldepth, t["order"].astype(np.int64)


def new_keys(x, y, w):
    return ob.Tree(ob.KIND_PIPE_QUAD, x, y, w).export()["keys"]


def repair(prev, ldepth, knew, shift, tile):
    """sort_repair_kernel + repair_movers_kernel, slot for slot.  Returns (kout, vout)."""
    n = len(prev)
    prev = [int(v) for v in prev]
    knew = [int(v) for v in knew]
    sh = [rp_shift(L) for L in ldepth]
    is_mover = [(knew[j] >> sh[j]) != (prev[j] >> sh[j]) for j in range(n)]
    movers = [j for j in range(n) if is_mover[j]]
    s_slot = sorted(movers)
    by_key = sorted(movers, key=lambda j: (knew[j] >> shift, j))
    t_sorted = [knew[j] >> shift for j in by_key]
    M = len(movers)

    def same_leaf(i, j):
        return 0 <= i < n and ldepth[i] == ldepth[j] and (prev[i] >> sh[i]) == (prev[j] >> sh[j])

    def cell_end(i):
        return ((prev[i] >> sh[i]) + 1) << sh[i]

    def lower(v, x):
        return bisect.bisect_left(v, x)

    kout, vout = [None] * n, [None] * n

    def put(p, key, slot):
        assert 0 <= p < n and kout[p] is None, ("two writers for position", p)
        kout[p], vout[p] = key, slot

    for j0 in range(0, n, tile):
        cnt = min(tile, n - j0)
        # keepers
        for j in range(j0, j0 + cnt):
            if is_mover[j]:
                continue
            t = knew[j] >> shift
            a, rank = j, 0
            while same_leaf(a - 1, j):
                assert j - (a - 1) < H, "leaf too long to rank in place (the kernel would fail the step)"
                ki = knew[a - 1]
                if (ki >> sh[j]) == (prev[j] >> sh[j]) and (ki >> shift) <= t:
                    rank += 1
                a -= 1
            i = j + 1
            while same_leaf(i, j):
                assert i - j < H
                ki = knew[i]
                if (ki >> sh[j]) == (prev[j] >> sh[j]) and (ki >> shift) < t:
                    rank += 1
                i += 1
            before = lower(s_slot, a)
            smaller = lower(t_sorted, t)
            put((a - before) + rank + smaller, knew[j], j)
        # the movers this tile places
        e_prev = 0 if j0 == 0 else cell_end(j0 - 1)
        e_last = cell_end(j0 + cnt - 1)
        up = (1 << shift) - 1
        r_lo = lower(t_sorted, (e_prev + up) >> shift)
        r_hi = M if j0 + cnt >= n else lower(t_sorted, (e_last + up) >> shift)
        for r in range(r_lo, r_hi):
            t = t_sorted[r]
            tk = t << shift
            lo = j0
            while lo < j0 + cnt and cell_end(lo) <= tk:
                lo += 1
            keepers = 0
            if lo < j0 + cnt and ((prev[lo] >> sh[lo]) << sh[lo]) <= tk:
                i = lo
                while same_leaf(i, lo):
                    assert i - lo < H
                    ki = knew[i]
                    if (ki >> sh[lo]) == (prev[lo] >> sh[lo]) and (ki >> shift) <= t:
                        keepers += 1
                    i += 1
            put((lo - lower(s_slot, lo)) + keepers + r, knew[by_key[r]], by_key[r])
    return kout, vout, M


def check(prev, ldepth, knew, shift, tile):
    kout, vout, M = repair(prev, ldepth, knew, shift, tile)
    n = len(prev)
    assert all(v is not None for v in vout), "a position nobody wrote"
    assert sorted(vout) == list(range(n)), "not a permutation of the slots"
    assert all(kout[p] == int(knew[vout[p]]) for p in range(n))
    t = [k >> shift for k in kout]
    assert all(t[p] <= t[p + 1] for p in range(n - 1)), "not sorted on the prefix"
    return M


def scene(n, seed):
    rng = np.random.default_rng(seed)
    # two clumps and a sparse background inside the default root square (0, 0, 100000): leaves of very different depth
    x = np.concatenate([rng.normal(30000, 1500, n // 2), rng.normal(70000, 300, n // 4), rng.uniform(5000, 95000, n - n // 2 - n // 4)])
    y = np.concatenate([rng.normal(40000, 1500, n // 2), rng.normal(60000, 300, n // 4), rng.uniform(5000, 95000, n - n // 2 - n // 4)])
    return x.astype(np.float32), y.astype(np.float32), np.ones(n, np.uint32), rng


@pytest.mark.parametrize("tile", [1024, 64, 7])
@pytest.mark.parametrize("motion", ["none", "jitter", "drift", "jumps", "to_the_ends"])
def test_repair_is_a_sort(tile, motion):
    n = 3000
    x, y, w, rng = scene(n, 11)
    prev, ldepth, order = old_tree(x, y, w)
    assert ldepth.max() <= 24 and ldepth.min() >= 1
    x2, y2 = x.copy(), y.copy()
    if motion == "jitter":      # inside the leaf cells, mostly: the keepers' in-leaf ranking
        x2 += rng.normal(0, 2.0, n).astype(np.float32)
        y2 += rng.normal(0, 2.0, n).astype(np.float32)
    elif motion == "drift":     # a few per cent cross cell lines
        x2 += rng.normal(0, 60.0, n).astype(np.float32)
        y2 += rng.normal(0, 60.0, n).astype(np.float32)
    elif motion == "jumps":     # far movers: into other leaves, into empty quadrants
        k = rng.choice(n, 150, replace=False)
        x2[k] = rng.uniform(100, 99900, len(k)).astype(np.float32)
        y2[k] = rng.uniform(100, 99900, len(k)).astype(np.float32)
    elif motion == "to_the_ends":  # before the first and behind the last body of the order, and all into one cell
        k = rng.choice(n, 90, replace=False)
        x2[k[:30]], y2[k[:30]] = 1.0, 1.0
        x2[k[30:60]], y2[k[30:60]] = 99999.0, 99999.0
        x2[k[60:]], y2[k[60:]] = 70000.0, 60000.0
    nk = new_keys(x2, y2, w)
    knew = nk[order]            # slot j holds body order[j]
    deepest = int(ldepth.max())
    for levels in sorted({deepest, deepest + 1, 24}):
        if levels > 24:
            continue
        shift = KEY_BITS - 2 * levels
        M = check(prev, ldepth, knew, shift, tile)
        if motion == "none":
            assert M == 0
        if motion in ("jumps", "to_the_ends"):
            assert M >= 80
