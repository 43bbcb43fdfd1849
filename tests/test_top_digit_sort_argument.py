"""The argument behind the quadtree build's top-digit sort (nbody_simulation_b200/csrc/radix_sort.cuh),
checked in numpy on the formulation tree_count_kernel uses:

  leaf depth L(i) = min(D, 1 + max over the windows of cap+1 consecutive sorted bodies containing i of the window's
  common prefix), and body i starts a leaf iff cpl(i-1, i) < L(i).

Claim: if the keys are stably sorted by their top S levels only, and no window of cap+1 consecutive bodies shares a
prefix of S levels or more, then leaf depths, leaf boundaries and leaf membership are exactly those of the full sort
(only the order inside a leaf differs, which tree_finish_kernel replaces by body-id order anyway).  And the flag
"some window shares >= S levels" is raised by the partially sorted keys whenever the fully sorted keys have a leaf
deeper than S.  No product code runs here: this pins the reasoning, the GPU tests pin the kernels."""
import numpy as np
import pytest

LEVELS = 24  # kQuadLevels: 48-bit keys, 2 bits per level


def cpl(a, b):
    """common prefix length in levels of 48-bit keys (elementwise)"""
    d = (a ^ b).astype(np.uint64)
    out = np.full(d.shape, LEVELS, dtype=np.int64)
    nz = d != 0
    # index of the highest set bit
    hb = np.zeros(d.shape, dtype=np.int64)
    v = d.copy()
    for sh in (32, 16, 8, 4, 2, 1):
        m = v >= (np.uint64(1) << np.uint64(sh))
        hb[m] += sh
        v[m] >>= np.uint64(sh)
    out[nz] = (47 - hb[nz]) // 2
    return out


def leaves(keys, cap):
    """(leaf depth per body, leaf-start flag per body, max window prefix) for keys in the given order"""
    n = len(keys)
    w = np.full(n + cap, -1, dtype=np.int64)  # w[t]: prefix of the window starting at body t - cap
    j = np.arange(-cap, n)
    ok = (j >= 0) & (j + cap < n)
    w[ok] = cpl(keys[j[ok]], keys[j[ok] + cap])
    cc = np.full(n, -1, dtype=np.int64)
    for q in range(cap + 1):  # windows starting at i - cap + q ... contain body i
        cc = np.maximum(cc, w[np.arange(n) + q])
    L = np.minimum(LEVELS, cc + 1)
    c = np.full(n, -1, dtype=np.int64)
    c[1:] = cpl(keys[:-1], keys[1:])
    return L, c < L, int(w.max())


def scene_keys(rng, n, clustered):
    if clustered:  # a dense clump: deep leaves
        x = np.concatenate([rng.random(n // 2), 0.3 + rng.random(n - n // 2) * 2.0 ** -9])
        y = np.concatenate([rng.random(n // 2), 0.7 + rng.random(n - n // 2) * 2.0 ** -9])
    else:
        x, y = rng.random(n), rng.random(n)
    xi = (x * (1 << LEVELS)).astype(np.uint64)
    yi = (y * (1 << LEVELS)).astype(np.uint64)
    k = np.zeros(n, dtype=np.uint64)
    for b in range(LEVELS):  # y bit above x bit (quad_tree.rs:179)
        k |= ((xi >> np.uint64(b)) & np.uint64(1)) << np.uint64(2 * b)
        k |= ((yi >> np.uint64(b)) & np.uint64(1)) << np.uint64(2 * b + 1)
    return k


@pytest.mark.parametrize("clustered", [False, True])
@pytest.mark.parametrize("cap", [1, 8, 33])
def test_top_level_sort_gives_the_same_leaves_or_raises_the_flag(clustered, cap):
    rng = np.random.default_rng(7 + cap)
    n = 6000
    keys = scene_keys(rng, n, clustered)
    ids = np.arange(n)
    full = np.argsort(keys, kind="stable")
    Lf, startf, _ = leaves(keys[full], cap)
    true_depth = int(Lf.max())
    leaf_of_full = np.cumsum(startf) - 1
    members_full = {}
    for pos, b in enumerate(ids[full]):
        members_full.setdefault(leaf_of_full[pos], []).append(b)
    canon_full = [sorted(v) for _, v in sorted(members_full.items())]
    checked_ok = checked_flag = 0
    for S in (4, 8, 12, 16, 20, 24):
        top = keys >> np.uint64(2 * (LEVELS - S))
        part = np.argsort(top, kind="stable")  # stable sort by the top S levels only
        Lp, startp, wmax = leaves(keys[part], cap)
        flag = wmax >= S
        if true_depth > S:
            assert flag, (S, true_depth, wmax)  # a leaf deeper than the sorted prefix is always noticed
            checked_flag += 1
            continue
        assert not flag or S == LEVELS, (S, true_depth, wmax)  # (at S = 24 the window prefix saturates at the depth cap)
        leaf_of = np.cumsum(startp) - 1
        members = {}
        for pos, b in enumerate(ids[part]):
            members.setdefault(leaf_of[pos], []).append(b)
        canon = [sorted(v) for _, v in sorted(members.items())]
        assert canon == canon_full, S  # same leaves, in the same order, holding the same bodies
        # per body: same leaf depth
        Lp_by_id = np.empty(n, dtype=np.int64)
        Lp_by_id[ids[part]] = Lp
        Lf_by_id = np.empty(n, dtype=np.int64)
        Lf_by_id[ids[full]] = Lf
        assert np.array_equal(Lp_by_id, Lf_by_id), S
        checked_ok += 1
    assert checked_ok >= 1 and (checked_flag >= 1 or true_depth <= 4)
