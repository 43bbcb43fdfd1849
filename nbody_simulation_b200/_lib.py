"""ctypes binding of libnbody_b200.so (include/nbody_b200.h).

The library is the product; this module only declares its C ABI.  There is no Python or CPU
fallback: if the shared object is missing, import fails loudly.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libnbody_b200.so")

NB_OK, NB_ERR_INVALID, NB_ERR_CUDA, NB_ERR_CAPACITY, NB_ERR_NCCL, NB_ERR_OVERFLOW = range(6)
NB_TREE_QUAD, NB_TREE_LBVH, NB_TREE_BVH_REF = 0, 1, 2
NB_FORCE_EXACT, NB_FORCE_FAST = 0, 1
NB_SCENE_DEFAULT, NB_SCENE_UNIFORM_DISC, NB_SCENE_PLUMMER, NB_SCENE_TWO_GALAXY = range(4)
NB_UNIQUE_ID_BYTES = 128
NB_PEER_BLOCK_BYTES = 640
NB_NOALLOC_MAX_TARGETS = 65536

# struct nb_particle (24 B) == reference `Particle` (src/main.rs:193-198)
PARTICLE_DTYPE = np.dtype(
    [("x", "<f4"), ("y", "<f4"), ("vx", "<f4"), ("vy", "<f4"), ("weight", "<u4"), ("_pad", "<u4")]
)
assert PARTICLE_DTYPE.itemsize == 24


class NbConfig(C.Structure):
    _fields_ = [
        ("n_max", C.c_uint64),
        ("theta", C.c_float),
        ("dt", C.c_float),
        ("tree_kind", C.c_int32),
        ("leaf_cap", C.c_int32),
        ("domain_x0", C.c_float),
        ("domain_y0", C.c_float),
        ("domain_h", C.c_float),
        ("domain_auto", C.c_int32),
        ("force_mode", C.c_int32),
        ("device", C.c_int32),
        ("faithful_index_bug", C.c_int32),
        ("full_sort", C.c_int32),
        ("fast_walk_group", C.c_int32),
        ("reserved", C.c_int32 * 6),
    ]


class NbKernelTime(C.Structure):
    _fields_ = [("name", C.c_char * 56), ("us", C.c_double)]


class NbCounting(C.Structure):
    _fields_ = [
        ("build_s", C.c_double),
        ("traverse_s", C.c_double),
        ("integrate_s", C.c_double),
        ("exchange_s", C.c_double),
        ("total_s", C.c_double),
        ("steps", C.c_uint64),
        ("interactions", C.c_uint64),
        ("node_tests", C.c_uint64),
        ("launches", C.c_uint64),
        ("reserved", C.c_uint64 * 3),
    ]


class NbTreeInfo(C.Structure):
    _fields_ = [
        ("n_bodies", C.c_uint64),
        ("n_in_tree", C.c_uint64),
        ("n_nodes", C.c_uint64),
        ("n_leaves", C.c_uint64),
        ("max_depth", C.c_uint32),
        ("overflow_leaves", C.c_uint32),
        ("root_x0", C.c_float),
        ("root_y0", C.c_float),
        ("root_h", C.c_float),
        ("sort_top_digits", C.c_uint32),
        ("sort_redone", C.c_uint32),
        ("sort_top_levels", C.c_uint32),
        ("sort_repaired", C.c_uint32),
        ("sort_movers", C.c_uint32),
    ]


TREE_VIEW_FIELDS = [
    ("keys", np.uint64, "n"),
    ("order", np.uint32, "n"),
    ("node_depth", np.int32, "nn"),
    ("node_start", np.uint32, "nn"),
    ("node_count", np.uint32, "nn"),
    ("node_skip", np.uint32, "nn"),
    ("node_parent", np.uint32, "nn"),
    ("node_leaf", np.uint8, "nn"),
    ("node_mass", np.uint32, "nn"),
    ("node_cogx", np.float32, "nn"),
    ("node_cogy", np.float32, "nn"),
    ("node_x0", np.float32, "nn"),
    ("node_y0", np.float32, "nn"),
    ("node_x1", np.float32, "nn"),
    ("node_y1", np.float32, "nn"),
    ("node_s2", np.float32, "nn"),
]


class NbTreeView(C.Structure):
    _fields_ = [(name, C.c_void_p) for name, _, _ in TREE_VIEW_FIELDS]


def alloc_tree_view(n, nn):
    """numpy buffers + the nb_tree_view pointing at them."""
    arrays = {}
    view = NbTreeView()
    for name, dt, which in TREE_VIEW_FIELDS:
        a = np.zeros(n if which == "n" else nn, dtype=dt)
        arrays[name] = a
        setattr(view, name, a.ctypes.data if a.size else None)
    return arrays, view


def load(path=LIB_PATH):
    if not os.path.exists(path):
        raise ImportError(
            f"{path} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(make -C nbody_simulation_b200/csrc). There is no CPU fallback."
        )
    lib = C.CDLL(path)
    P = C.POINTER
    vp, f32p, u32p = C.c_void_p, C.c_void_p, C.c_void_p
    sigs = {
        "nb_abi_version": (C.c_int, []),
        "nb_default_config": (None, [P(NbConfig), C.c_uint64]),
        "nb_create": (vp, [P(NbConfig)]),
        "nb_destroy": (None, [vp]),
        "nb_last_error": (C.c_char_p, [vp]),
        "nb_set_params": (C.c_int, [vp, C.c_float, C.c_float, C.c_int32]),
        "nb_set_counting": (C.c_int, [vp, C.c_int]),
        "nb_set_bodies_aos": (C.c_int, [vp, vp, C.c_size_t]),
        "nb_set_bodies_soa": (C.c_int, [vp, f32p, f32p, f32p, f32p, u32p, C.c_size_t]),
        "nb_get_bodies_aos": (C.c_int, [vp, vp, C.c_size_t]),
        "nb_get_bodies_soa": (C.c_int, [vp, f32p, f32p, f32p, f32p, u32p, C.c_size_t]),
        "nb_step": (C.c_int, [vp, C.c_uint32, P(NbCounting)]),
        "nb_world_update": (C.c_int, [vp, vp, C.c_size_t, C.c_float, P(NbCounting)]),
        "nb_build_tree": (C.c_int, [vp, P(NbCounting)]),
        "nb_accel": (C.c_int, [vp, f32p, f32p, P(NbCounting)]),
        "nb_accel_direct": (C.c_int, [vp, C.c_size_t, C.c_size_t, f32p, f32p, P(C.c_double)]),
        "nb_host_alloc": (vp, [C.c_size_t]),
        "nb_host_free": (None, [vp]),
        "nb_flush_l2": (C.c_int, [vp]),
        "nb_measure_fp32": (C.c_int, [vp, C.c_int, P(C.c_double)]),
        "nb_get_tree_info": (C.c_int, [vp, P(NbTreeInfo)]),
        "nb_get_tree": (C.c_int, [vp, P(NbTreeView)]),
        "nb_draw": (C.c_int, [vp, vp, C.c_uint32]),
        "nb_world_update_shard": (C.c_int, [vp, vp, C.c_size_t, C.c_size_t, C.c_size_t, C.c_float, vp]),
        "nb_quad_empty": (C.c_int, [vp, C.POINTER(C.c_uint32)]),
        "nb_quad_prune": (C.c_int, [vp, C.POINTER(C.c_uint32)]),
        "nb_body_count": (C.c_size_t, [vp]),
        "nb_accel_direct_noalloc": (C.c_int, [vp, vp, vp, C.c_size_t, C.c_size_t, C.c_size_t, C.c_int32, vp, vp, vp]),
        "nb_profile_step": (C.c_int, [vp, vp, C.c_size_t, C.POINTER(C.c_size_t)]),
        "nb_save_snapshot": (C.c_int, [vp, C.c_char_p, C.c_uint64]),
        "nb_load_snapshot": (C.c_int, [vp, C.c_char_p, P(C.c_uint64)]),
        "nb_comm_unique_id": (C.c_int, [vp]),
        "nb_comm_init": (C.c_int, [vp, C.c_int, C.c_int, vp]),
        "nb_comm_owned_range": (C.c_int, [vp, P(C.c_uint64), P(C.c_uint64)]),
        "nb_comm_peer_export": (C.c_int, [vp, vp]),
        "nb_comm_peer_import": (C.c_int, [vp, vp]),
        "nb_comm_peer_enable": (C.c_int, [vp, C.c_int]),
        "nb_shard_range": (None, [C.c_uint64, C.c_int, C.c_int, P(C.c_uint64), P(C.c_uint64)]),
        "nb_scene_generate": (C.c_size_t, [C.c_int32, C.c_uint64, C.c_size_t, vp, C.c_size_t]),
        "nb_scene_generate_device": (C.c_int, [vp, C.c_int32, C.c_uint64, C.c_size_t]),
    }
    for name, (res, args) in sigs.items():
        fn = getattr(lib, name)  # AttributeError if the .so does not export what the header declares
        fn.restype = res
        fn.argtypes = args
    lib._nb_symbols = tuple(sigs)
    return lib


_lib = None


def lib():
    global _lib
    if _lib is None:
        # NB_B200_LIB: another build of the same library (A/B runs of a kernel change: scripts/ab_step.py)
        _lib = load(os.environ.get("NB_B200_LIB") or LIB_PATH)
    return _lib
