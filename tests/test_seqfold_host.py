"""CPU tier: the exact parallel formulation of the sequential f32 fold (csrc/seqfold_core.h, the
split-point sum of BVHTree::from, reference src/bvh_tree.rs:58-67) equals a plain left-to-right
f32 loop bit for bit, on positions like the scenes' and on adversarial inputs."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def shim():
    out = os.path.join(HERE, "_build")
    os.makedirs(out, exist_ok=True)
    so = os.path.join(out, "libseqfold_shim.so")
    subprocess.check_call(["g++", "-O2", "-ffp-contract=off", "-fno-fast-math", "-fPIC", "-shared", "-std=c++17",
                           "-o", so, os.path.join(HERE, "hostshim", "seqfold_shim.cpp")])
    L = ctypes.CDLL(so)
    fp = ctypes.POINTER(ctypes.c_float)
    L.nbt_fold_plain.restype = ctypes.c_float
    L.nbt_fold_plain.argtypes = [fp, ctypes.c_size_t, ctypes.c_float]
    L.nbt_fold_windowed.restype = ctypes.c_float
    L.nbt_fold_windowed.argtypes = [fp, ctypes.c_size_t, ctypes.c_float, ctypes.c_int, ctypes.c_int,
                                    ctypes.POINTER(ctypes.c_ulonglong)]
    L.nbt_fold_tiles.restype = ctypes.c_float
    L.nbt_fold_tiles.argtypes = [fp, ctypes.c_size_t, ctypes.c_int, ctypes.POINTER(ctypes.c_ulonglong)]
    L.nbt_fold_scan.restype = ctypes.c_float
    L.nbt_fold_scan.argtypes = [fp, ctypes.c_size_t, ctypes.c_float, ctypes.c_int, ctypes.c_int]
    return L


def both(L, x, acc=0.0, chunk=8, threads=64):
    x = np.ascontiguousarray(x, dtype=np.float32)
    p = x.ctypes.data_as(ctypes.POINTER(ctypes.c_float))
    slow = ctypes.c_ulonglong(0)
    a = np.float32(L.nbt_fold_plain(p, len(x), acc))
    b = np.float32(L.nbt_fold_windowed(p, len(x), acc, chunk, threads, ctypes.byref(slow)))
    if acc == 0.0:  # the tile pipeline of the big-node fold (fp64 binade guesses, plain fallback), tiles of 256
        d = np.float32(L.nbt_fold_tiles(p, len(x), 256, None))
        assert a.view(np.uint32) == d.view(np.uint32) or (np.isnan(a) and np.isnan(d)), "tile pipeline differs"
    c = np.float32(L.nbt_fold_scan(p, len(x), acc, chunk, threads))
    assert b.view(np.uint32) == c.view(np.uint32) or (np.isnan(b) and np.isnan(c)), "scan-shaped variant differs"
    return a, b, slow.value


def same(a, b):
    return a.view(np.uint32) == b.view(np.uint32) or (np.isnan(a) and np.isnan(b))


CASES = {
    "positions": lambda r, n: r.uniform(0, 1e5, n),
    "disc": lambda r, n: 5e4 + 4.5e4 * np.sqrt(r.uniform(0, 1, n)) * np.cos(r.uniform(0, 6.28, n)),
    "mixed_sign": lambda r, n: r.normal(0, 1e3, n),
    "ties": lambda r, n: r.integers(0, 1 << 12, n) * 0.5,
    "tiny_and_huge": lambda r, n: r.choice([1e-30, 3.5, 1e20, -1e20, 7e-3, 1e5], n),
    "shrinking": lambda r, n: np.concatenate([[1e9], -r.uniform(0, 2e4, max(n - 1, 0))]),
    "subnormal": lambda r, n: r.uniform(0, 1e-40, n),
    "constant": lambda r, n: np.full(n, 12345.678),
    "same_binade_runs": lambda r, n: np.tile(np.array([3e4, 3.1e4, -6.05e4, 2.9e4, 3.3e4], dtype=np.float32), n // 5 + 1)[:n],
    "with_inf": lambda r, n: np.concatenate([r.uniform(0, 1e5, max(n - 2, 0)), [np.inf, 3.0]]),
}


@pytest.mark.parametrize("name", sorted(CASES))
@pytest.mark.parametrize("n", [1, 7, 1000, 200_000])
def test_windowed_fold_is_bit_exact(shim, name, n):
    r = np.random.default_rng(sorted(CASES).index(name) * 1000003 + n)
    x = np.asarray(CASES[name](r, n), dtype=np.float32)
    for chunk, threads in ((8, 64), (3, 5), (32, 256)):  # (32, 256) is the device kernel's window
        a, b, _ = both(shim, x, 0.0, chunk, threads)
        assert same(a, b), (name, n, chunk, threads, a, b)


def test_fast_path_covers_almost_everything(shim):
    # positions in the reference's domain: only the ~25 binade crossings take a plain step
    r = np.random.default_rng(5)
    x = r.uniform(0, 1e5, 1_000_000).astype(np.float32)
    a, b, slow = both(shim, x, 0.0, 8, 1024)
    assert same(a, b) and slow < 64, slow


def test_nonzero_start_and_negative_accumulator(shim):
    r = np.random.default_rng(6)
    x = r.uniform(-10, 1e3, 50_000).astype(np.float32)
    for acc in (-1e7, 3.25, 1e-3, -0.0):
        a, b, _ = both(shim, x, acc)
        assert same(a, b), acc


def test_tile_pipeline_rarely_falls_back(shim):
    # one plain tile per power of two the sum crosses, plus tile 0
    r = np.random.default_rng(7)
    x = r.uniform(0, 1e5, 1_000_000).astype(np.float32)
    plain = ctypes.c_ulonglong(0)
    v = np.float32(shim.nbt_fold_tiles(x.ctypes.data_as(ctypes.POINTER(ctypes.c_float)), len(x), 256, ctypes.byref(plain)))
    ref = np.float32(shim.nbt_fold_plain(x.ctypes.data_as(ctypes.POINTER(ctypes.c_float)), len(x), 0.0))
    assert same(v, ref) and plain.value <= 20, plain.value
