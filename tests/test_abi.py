"""The C-ABI library loads without a GPU and exports every symbol include/nbody_b200.h declares.
No compute calls here (no GPU in this tier); creating a context must fail LOUDLY, not fall back."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import nbody_simulation_b200 as nb
from nbody_simulation_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_functions():
    src = open(os.path.join(ROOT, "include", "nbody_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(nb_[a-z0-9_]+)\s*\(", src)))


def test_every_declared_symbol_is_exported_and_bound():
    L = _lib.lib()
    names = header_functions()
    assert len(names) >= 25
    for n in names:
        assert hasattr(L, n), f"{n} declared in the header but not exported"
    assert set(names) == set(L._nb_symbols), set(names) ^ set(L._nb_symbols)


def test_abi_version_and_struct_sizes():
    L = _lib.lib()
    assert L.nb_abi_version() == 1
    assert _lib.PARTICLE_DTYPE.itemsize == 24  # repr(C) Particle, src/main.rs:193-198
    assert C.sizeof(_lib.NbConfig) == 88  # 84 bytes of fields, 8-byte aligned
    assert C.sizeof(_lib.NbCounting) == 5 * 8 + 7 * 8
    assert C.sizeof(_lib.NbTreeInfo) == 4 * 8 + 2 * 4 + 3 * 4 + 5 * 4
    # fields carved out of the reserved words (ABI version unchanged: old callers passed / read zeros there)
    assert _lib.NbConfig.full_sort.offset == 52 and _lib.NbConfig.fast_walk_group.offset == 56 and _lib.NbConfig.reserved.offset == 60
    assert _lib.NbTreeInfo.sort_top_digits.offset == 52 and _lib.NbTreeInfo.sort_redone.offset == 56
    assert _lib.NbTreeInfo.sort_top_levels.offset == 60 and _lib.NbTreeInfo.sort_repaired.offset == 64 and _lib.NbTreeInfo.sort_movers.offset == 68
    cfg = _lib.NbConfig()
    L.nb_default_config(C.byref(cfg), 10)
    assert cfg.full_sort == 0 and not any(cfg.reserved)


def test_default_config_is_the_reference_constants():
    cfg = _lib.NbConfig()
    _lib.lib().nb_default_config(C.byref(cfg), 1234)
    assert cfg.n_max == 1234 and cfg.theta == 50.0 and abs(cfg.dt - 0.1) < 1e-7  # main.rs:34-35
    assert cfg.domain_h == 100000.0 and cfg.tree_kind == nb.NB_TREE_QUAD  # main.rs:31


def _no_gpu():
    try:
        import torch
        return not torch.cuda.is_available()
    except Exception:
        return True


@pytest.mark.skipif(not _no_gpu(), reason="only meaningful without a CUDA device")
def test_no_cpu_fallback():
    with pytest.raises(nb.NbError) as e:
        nb.World(1000)
    assert "no CPU fallback" in str(e.value) or "CUDA" in str(e.value)


def test_invalid_config_is_rejected():
    L = _lib.lib()
    cfg = _lib.NbConfig()
    L.nb_default_config(C.byref(cfg), 0)
    assert not L.nb_create(C.byref(cfg))
    assert b"n_max" in L.nb_last_error(None)
    L.nb_default_config(C.byref(cfg), 10)
    cfg.tree_kind = 7
    assert not L.nb_create(C.byref(cfg))
    cfg.tree_kind = 0
    cfg.leaf_cap = 65
    assert not L.nb_create(C.byref(cfg))


def test_scene_generator_is_seeded_and_matches_world_new():
    a = nb.generate_scene(nb.NB_SCENE_DEFAULT, 1)
    b = nb.generate_scene(nb.NB_SCENE_DEFAULT, 1)
    c = nb.generate_scene(nb.NB_SCENE_DEFAULT, 2)
    assert np.array_equal(a.view(np.uint8), b.view(np.uint8)) and len(a) != 0
    assert not np.array_equal(a[:len(c)].view(np.uint8), c[:len(a)].view(np.uint8))
    # main.rs:282-291: the two heavy bodies
    assert tuple(a[0])[:5] == (35000.0, 35000.0, 200.0, 250.0, 75_000_000)
    assert tuple(a[1])[:5] == (60000.0, 60000.0, 0.0, 0.0, 750_000)
    # main.rs:334-335: exactly 100000 rand_body after the grid disc; all weight 1
    assert len(a) > 100_002 and (a["weight"][2:] == 1).all()
    tail = a[-100_000:]
    r = np.hypot(tail["x"] - 50000.0, tail["y"] - 50000.0)
    assert r.max() <= 25000.0 * 1.0001 and np.hypot(tail["vx"], tail["vy"]).max() <= 1.0001
    grid = a[2:-100_000]  # main.rs:316-332: multiples of 14 inside the annulus, |v| = 750000^(1/4)
    assert (np.mod(grid["x"], 14.0) == 0).all() and (np.mod(grid["y"], 14.0) == 0).all()
    d2 = (grid["x"] - 60000.0) ** 2 + (grid["y"] - 60000.0) ** 2
    assert d2.min() > 5e5 and d2.max() < 9e6 + 1
    assert np.allclose(np.hypot(grid["vx"], grid["vy"]), 750000.0 ** 0.25, rtol=1e-4)
    for kind in (nb.NB_SCENE_UNIFORM_DISC, nb.NB_SCENE_PLUMMER, nb.NB_SCENE_TWO_GALAXY):
        s = nb.generate_scene(kind, 5, 10_001)
        assert len(s) == 10_001
        assert (s["x"] >= 0).all() and (s["x"] < 100000).all() and (s["y"] >= 0).all() and (s["y"] < 100000).all()


def test_shard_rule():
    for n in (0, 1, 7, 1000, 1_000_003):
        for g in (1, 2, 4, 8):
            spans = [nb.shard_range(n, r, g) for r in range(g)]
            assert sum(c for _, c in spans) == n
            pos = 0
            for f, c in spans:
                assert f == pos or c == 0
                pos += c
            per = -(-(-(-n // g)) // 32) * 32 if g > 1 else n  # ceil(n / g) rounded up to whole 32-body warps
            assert all(c <= per for _, c in spans)
            assert all(f % 32 == 0 for f, c in spans if c)
