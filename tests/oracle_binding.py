"""ctypes binding of the CPU oracle (oracle/nbody_oracle.cpp).  TEST INFRASTRUCTURE ONLY."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
ORACLE_SO = os.path.join(ORACLE_DIR, "_build", "libnbody_oracle.so")
NATIVE_SO = os.path.join(ORACLE_DIR, "_build", "libnbody_oracle_native.so")
NATIVE_FLAGS = "-O3 -march=native -ffp-contract=off -fno-fast-math -fopenmp -fPIC -std=c++17 -Wall -Wno-unknown-pragmas"
PORTABLE_FLAGS = "-O3 -march=x86-64-v3 -ffp-contract=off -fno-fast-math -fopenmp -fPIC -std=c++17"
build_flags = PORTABLE_FLAGS  # of the library lib() loaded

# struct layouts only: importing the package does not load libnbody_b200.so (that happens on the first call)
from nbody_simulation_b200._lib import (PARTICLE_DTYPE, NbCounting, NbTreeInfo, NbTreeView,  # noqa: E402
                                        alloc_tree_view)

KIND_PIPE_QUAD, KIND_PIPE_LBVH, KIND_LITERAL_QUAD, KIND_LITERAL_BVH = 0, 1, 2, 3

_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    # `make` is a no-op when the library is newer than its source and the shared header
    global build_flags
    so = ORACLE_SO
    if os.environ.get("NB_ORACLE_NATIVE") == "1":
        # timed CPU baseline: built on THIS machine for its own cores (-O3 -march=native)
        try:
            subprocess.check_call(["make", "-C", ORACLE_DIR, f"OUT={NATIVE_SO}", f"CXXFLAGS={NATIVE_FLAGS}"],
                                  stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL, timeout=300)
            so, build_flags = NATIVE_SO, NATIVE_FLAGS
        except Exception:
            so = ORACLE_SO
    if so == ORACLE_SO:
        subprocess.check_call(["make", "-C", ORACLE_DIR], stdout=subprocess.DEVNULL)
    L = C.CDLL(so)
    vp, f, P = C.c_void_p, C.c_float, C.POINTER
    L.orc_pair_force.argtypes = [f, f, f, f, f, P(f), P(f)]
    L.orc_pair_force.restype = None
    L.orc_integrate.argtypes = [vp, vp, vp, vp, vp, vp, C.c_size_t, f]
    L.orc_integrate.restype = None
    L.orc_opening_accept.argtypes = [f] * 9
    L.orc_opening_accept.restype = C.c_int
    L.orc_morton_key.argtypes = [f] * 5
    L.orc_morton_key.restype = C.c_uint64
    L.orc_tree_build.argtypes = [C.c_int, vp, vp, vp, C.c_size_t, f, f, f, C.c_int, C.c_int]
    L.orc_tree_build.restype = vp
    L.orc_tree_free.argtypes = [vp]
    L.orc_tree_free.restype = None
    L.orc_tree_info.argtypes = [vp, P(NbTreeInfo)]
    L.orc_tree_export.argtypes = [vp, P(NbTreeView)]
    L.orc_tree_accel.argtypes = [vp, f, vp, C.c_size_t, C.c_int, vp, vp, P(C.c_uint64), P(C.c_uint64), C.c_int,
                                 P(C.c_double)]
    L.orc_quad_empty_prune.argtypes = [vp, P(C.c_uint32), P(C.c_uint32)]
    L.orc_quad_fresh_stats.argtypes = [vp, P(C.c_uint32), P(C.c_uint32)]
    L.orc_direct_f64.argtypes = [vp, vp, vp, C.c_size_t, vp, C.c_size_t, vp, vp, C.c_int]
    L.orc_direct_f32.argtypes = [vp, vp, vp, C.c_size_t, vp, C.c_size_t, vp, vp, C.c_int]
    L.orc_world_new.argtypes = [vp, vp, vp, vp, vp, C.c_size_t]
    L.orc_world_new.restype = vp
    L.orc_world_free.argtypes = [vp]
    L.orc_world_free.restype = None
    L.orc_world_get.argtypes = [vp, vp, vp, vp, vp, vp]
    L.orc_world_get.restype = None
    L.orc_world_update.argtypes = [vp, C.c_int, f, f, f, f, f, C.c_int, C.c_int, C.c_int, C.c_size_t, P(NbCounting)]
    L.orc_num_threads.restype = C.c_int
    L.nb_scene_generate.argtypes = [C.c_int32, C.c_uint64, C.c_size_t, vp, C.c_size_t]
    L.nb_scene_generate.restype = C.c_size_t
    _lib = L
    return L


def generate_scene(scene, seed=1, n=0):
    """The seeded scenes of csrc/scene.cpp as compiled into the ORACLE library (same source file, same bits as
    nbody_simulation_b200.generate_scene): inputs for the reference arm without loading the product library."""
    L = lib()
    total = L.nb_scene_generate(scene, seed, n, None, 0)
    out = np.zeros(total, dtype=PARTICLE_DTYPE)
    assert L.nb_scene_generate(scene, seed, n, out.ctypes.data, total) == total
    return out


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _u32(a):
    return np.ascontiguousarray(a, dtype=np.uint32)


class Tree:
    def __init__(self, kind, x, y, w, domain=(0.0, 0.0, 100000.0), cap=0, domain_auto=False):
        self.L = lib()
        self.x, self.y, self.w = _f32(x), _f32(y), _u32(w)
        self.n = len(self.x)
        self.kind = kind
        self.h = self.L.orc_tree_build(kind, self.x.ctypes.data, self.y.ctypes.data, self.w.ctypes.data, self.n,
                                       domain[0], domain[1], domain[2], cap, 1 if domain_auto else 0)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orc_tree_free(self.h)
            self.h = None

    def info(self):
        info = NbTreeInfo()
        self.L.orc_tree_info(self.h, C.byref(info))
        return info

    def export(self):
        info = self.info()
        arrays, view = alloc_tree_view(int(info.n_bodies), int(info.n_nodes))
        self.L.orc_tree_export(self.h, C.byref(view))
        arrays["info"] = info
        return arrays

    def accel(self, theta, targets=None, literal=False, threads=0):
        if targets is None:
            nt, tp = self.n, None
        else:
            targets = _u32(targets)
            nt, tp = len(targets), targets.ctypes.data
        ax, ay = np.zeros(nt, np.float32), np.zeros(nt, np.float32)
        ni, ntst, sec = C.c_uint64(), C.c_uint64(), C.c_double()
        self.L.orc_tree_accel(self.h, theta, tp, nt, 1 if literal else 0, ax.ctypes.data, ay.ctypes.data,
                              C.byref(ni), C.byref(ntst), threads, C.byref(sec))
        return ax, ay, ni.value, ntst.value, sec.value

    def fresh_stats(self):
        """(empty() return value, prune() return value) of the fresh literal quadtree; prune mutates the tree."""
        e, p = C.c_uint32(), C.c_uint32()
        assert self.L.orc_quad_fresh_stats(self.h, C.byref(e), C.byref(p)) == 0
        return e.value, p.value

    def empty_prune(self):
        e, p = C.c_uint32(), C.c_uint32()
        rc = self.L.orc_quad_empty_prune(self.h, C.byref(e), C.byref(p))
        assert rc == 0
        return e.value, p.value


def direct_f64(x, y, w, targets=None, threads=0):
    L = lib()
    x, y, w = _f32(x), _f32(y), _u32(w)
    if targets is None:
        nt, tp = len(x), None
    else:
        targets = _u32(targets)
        nt, tp = len(targets), targets.ctypes.data
    ax, ay = np.zeros(nt, np.float64), np.zeros(nt, np.float64)
    L.orc_direct_f64(x.ctypes.data, y.ctypes.data, w.ctypes.data, len(x), tp, nt, ax.ctypes.data, ay.ctypes.data, threads)
    return ax, ay


def direct_f32(x, y, w, targets=None, threads=0):
    L = lib()
    x, y, w = _f32(x), _f32(y), _u32(w)
    if targets is None:
        nt, tp = len(x), None
    else:
        targets = _u32(targets)
        nt, tp = len(targets), targets.ctypes.data
    ax, ay = np.zeros(nt, np.float32), np.zeros(nt, np.float32)
    L.orc_direct_f32(x.ctypes.data, y.ctypes.data, w.ctypes.data, len(x), tp, nt, ax.ctypes.data, ay.ctypes.data, threads)
    return ax, ay


class OracleWorld:
    """World + World::update of the reference on the CPU oracle."""

    def __init__(self, parts):
        self.L = lib()
        self.n = len(parts)
        a = [_f32(parts[k]) for k in ("x", "y", "vx", "vy")]
        w = _u32(parts["weight"])
        self.h = self.L.orc_world_new(*[v.ctypes.data for v in a], w.ctypes.data, self.n)

    def __del__(self):
        if getattr(self, "h", None):
            self.L.orc_world_free(self.h)
            self.h = None

    def update(self, kind, delta, theta, domain=(0.0, 0.0, 100000.0), cap=0, faithful_bug=False, threads=0,
               sample_stride=1, counting=None):
        c = counting if counting is not None else NbCounting()
        rc = self.L.orc_world_update(self.h, kind, delta, theta, domain[0], domain[1], domain[2], cap,
                                     1 if faithful_bug else 0, threads, sample_stride, C.byref(c))
        assert rc == 0
        return c

    def get(self):
        x, y, vx, vy = (np.zeros(self.n, np.float32) for _ in range(4))
        w = np.zeros(self.n, np.uint32)
        self.L.orc_world_get(self.h, x.ctypes.data, y.ctypes.data, vx.ctypes.data, vy.ctypes.data, w.ctypes.data)
        return x, y, vx, vy, w


def normwise_err(ax, ay, rx, ry):
    """||da||_2 / ||a||_2 over all bodies (the tolerance norm used throughout)."""
    ax, ay, rx, ry = (np.asarray(v, np.float64) for v in (ax, ay, rx, ry))
    num = np.sqrt(np.sum((ax - rx) ** 2 + (ay - ry) ** 2))
    den = np.sqrt(np.sum(rx ** 2 + ry ** 2))
    return float(num / den) if den > 0 else float(num)
