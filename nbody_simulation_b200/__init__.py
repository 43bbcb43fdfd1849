"""nbody_simulation_b200 — host-side mirror of the reference's simulation API over the
B200-native C-ABI library (libnbody_b200.so, include/nbody_b200.h).

Reference surface mirrored (KristinnVikarJ/nbody-simulation, all in src/main.rs):
  World { particles }            :37-39   -> World.particles
  World::new()                   :276-346 -> World.new(seed=...)  (seeded restatement)
  World::update(delta, counter)  :388-425 -> World.update(delta, counter)
  Counting { build_bvh, sum_gravity, post_calculations } :74-79 -> Counting
  consts THETA / STEP_SIZE / HEIGHT :31-35 -> THETA / STEP_SIZE / HEIGHT defaults
  draw(particles, frame)         :41-72   -> World.draw()
Compute never happens here: every method is one call through the C ABI.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import (NB_FORCE_EXACT, NB_FORCE_FAST, NB_SCENE_DEFAULT, NB_SCENE_PLUMMER,
                   NB_SCENE_TWO_GALAXY, NB_SCENE_UNIFORM_DISC, NB_TREE_BVH_REF, NB_TREE_LBVH, NB_TREE_QUAD,
                   PARTICLE_DTYPE, NbConfig, NbCounting, NbTreeInfo)

HEIGHT = 100_000        # src/main.rs:31
RENDER_HEIGHT = 1250    # src/main.rs:32
STEP_SIZE = 0.1         # src/main.rs:34
THETA = 50.0            # src/main.rs:35

_STATUS = {1: "invalid argument/state", 2: "CUDA error", 3: "capacity", 4: "NCCL error", 5: "node pool overflow"}


class NbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"nbody_b200 error {code} ({_STATUS.get(code, '?')}): {msg}")
        self.code = code


class Counting:
    """`struct Counting` (src/main.rs:74-79): cumulative seconds per phase."""

    def __init__(self):
        self.c = NbCounting()

    build_bvh = property(lambda s: s.c.build_s)
    sum_gravity = property(lambda s: s.c.traverse_s)
    post_calculations = property(lambda s: s.c.integrate_s)
    exchange = property(lambda s: s.c.exchange_s)
    total = property(lambda s: s.c.total_s)
    steps = property(lambda s: s.c.steps)
    interactions = property(lambda s: s.c.interactions)
    node_tests = property(lambda s: s.c.node_tests)
    launches = property(lambda s: s.c.launches)

    def __repr__(self):  # the reference prints `{:?}` once per second (main.rs:152)
        return (f"Counting {{ build_bvh: {self.build_bvh}, sum_gravity: {self.sum_gravity}, "
                f"post_calculations: {self.post_calculations} }}")


def generate_scene(scene, seed=1, n=0):
    """Seeded synthetic scenes (host code inside the library, no GPU needed)."""
    L = _lib.lib()
    total = L.nb_scene_generate(scene, seed, n, None, 0)
    out = np.zeros(total, dtype=PARTICLE_DTYPE)
    got = L.nb_scene_generate(scene, seed, n, out.ctypes.data, total)
    assert got == total
    return out


def shard_range(n, rank, n_ranks):
    """Morton-range sharding rule of the multi-GPU step (host arithmetic inside the library)."""
    f, c = C.c_uint64(), C.c_uint64()
    _lib.lib().nb_shard_range(int(n), int(rank), int(n_ranks), C.byref(f), C.byref(c))
    return f.value, c.value


def pinned_particles(n):
    """A particle array in pinned host memory (nb_host_alloc); keep the returned owner alive."""
    L = _lib.lib()
    nbytes = max(1, n) * PARTICLE_DTYPE.itemsize
    p = L.nb_host_alloc(nbytes)
    if not p:
        raise NbError(2, "nb_host_alloc failed")
    buf = (C.c_char * nbytes).from_address(p)
    arr = np.frombuffer(buf, dtype=PARTICLE_DTYPE, count=n)

    class _Owner:
        def __init__(self, ptr):
            self.ptr = ptr

        def __del__(self):
            try:
                L.nb_host_free(self.ptr)
            except Exception:
                pass

    return arr, _Owner(p)


class World:
    """`struct World` + `impl World` of the reference, state resident in HBM."""

    def __init__(self, n_max, theta=THETA, dt=STEP_SIZE, tree=NB_TREE_QUAD, leaf_cap=0,
                 domain=(0.0, 0.0, float(HEIGHT)), domain_auto=False, force_mode=NB_FORCE_FAST, device=0,
                 faithful_index_bug=False, full_sort=False, fast_walk_group=False):
        self._L = _lib.lib()
        cfg = NbConfig()
        self._L.nb_default_config(C.byref(cfg), int(n_max))
        cfg.theta, cfg.dt = float(theta), float(dt)
        cfg.tree_kind, cfg.leaf_cap = int(tree), int(leaf_cap)
        cfg.domain_x0, cfg.domain_y0, cfg.domain_h = map(float, domain)
        cfg.domain_auto = 1 if domain_auto else 0
        cfg.force_mode, cfg.device = int(force_mode), int(device)
        cfg.faithful_index_bug = 1 if faithful_index_bug else 0
        cfg.full_sort = 1 if full_sort else 0
        cfg.fast_walk_group = int(fast_walk_group)
        self.cfg = cfg
        self._ctx = self._L.nb_create(C.byref(cfg))
        if not self._ctx:
            raise NbError(2, self._L.nb_last_error(None).decode())
        self.n = 0

    @classmethod
    def new(cls, seed=1, **kw):
        """`World::new()` (src/main.rs:276-346), seeded."""
        parts = generate_scene(NB_SCENE_DEFAULT, seed)
        w = cls(len(parts), **kw)
        w.set_particles(parts)
        return w

    def close(self):
        if getattr(self, "_ctx", None):
            self._L.nb_destroy(self._ctx)
            self._ctx = None

    __del__ = close

    def _ck(self, rc):
        if rc != 0:
            raise NbError(rc, self._L.nb_last_error(self._ctx).decode())

    # ---- state -------------------------------------------------------------------------
    def set_particles(self, parts):
        parts = np.ascontiguousarray(parts, dtype=PARTICLE_DTYPE)
        self._ck(self._L.nb_set_bodies_aos(self._ctx, parts.ctypes.data, len(parts)))
        self.n = len(parts)

    def set_soa(self, x, y, vx, vy, w):
        a = [np.ascontiguousarray(v, dtype=np.float32) for v in (x, y, vx, vy)]
        w = np.ascontiguousarray(w, dtype=np.uint32)
        self._ck(self._L.nb_set_bodies_soa(self._ctx, *[v.ctypes.data for v in a], w.ctypes.data, len(w)))
        self.n = len(w)

    @property
    def particles(self):
        out = np.zeros(self.n, dtype=PARTICLE_DTYPE)
        self._ck(self._L.nb_get_bodies_aos(self._ctx, out.ctypes.data, self.n))
        return out

    def get_soa(self):
        x, y, vx, vy = (np.zeros(self.n, np.float32) for _ in range(4))
        w = np.zeros(self.n, np.uint32)
        self._ck(self._L.nb_get_bodies_soa(self._ctx, x.ctypes.data, y.ctypes.data, vx.ctypes.data,
                                           vy.ctypes.data, w.ctypes.data, self.n))
        return x, y, vx, vy, w

    def set_params(self, theta=None, dt=None, force_mode=None):
        if theta is not None:
            self.cfg.theta = float(theta)
        if dt is not None:
            self.cfg.dt = float(dt)
        if force_mode is not None:
            self.cfg.force_mode = int(force_mode)
        self._ck(self._L.nb_set_params(self._ctx, self.cfg.theta, self.cfg.dt, self.cfg.force_mode))

    def set_counting(self, enabled):
        self._ck(self._L.nb_set_counting(self._ctx, 1 if enabled else 0))

    # ---- the step ------------------------------------------------------------------------
    def update(self, delta=None, counter=None):
        """`World::update(delta, &mut counter)` (src/main.rs:388-425)."""
        if delta is not None and float(delta) != self.cfg.dt:
            self.set_params(dt=delta)
        self.step(1, counter)

    def step(self, n_steps=1, counter=None):
        self._ck(self._L.nb_step(self._ctx, int(n_steps), C.byref(counter.c) if counter else None))

    def update_host(self, parts, delta=None, counter=None):
        """Drop-in form with the particle Vec on the host: upload, one step, download in place."""
        assert parts.dtype == PARTICLE_DTYPE and parts.flags["C_CONTIGUOUS"]
        d = self.cfg.dt if delta is None else float(delta)
        self._ck(self._L.nb_world_update(self._ctx, parts.ctypes.data, len(parts), d,
                                         C.byref(counter.c) if counter else None))
        self.n = len(parts)

    def update_host_shard(self, slice_, first, n_total, delta=None, counter=None):
        """World::update for a Vec sharded over the ranks' hosts (collective): this rank holds bodies
        [first, first + len(slice_)) of the n_total-body Vec; upload 1/G, NVLink to all, step, download 1/G."""
        assert slice_.dtype == PARTICLE_DTYPE and slice_.flags["C_CONTIGUOUS"]
        d = self.cfg.dt if delta is None else float(delta)
        self._ck(self._L.nb_world_update_shard(self._ctx, slice_.ctypes.data, int(first), len(slice_), int(n_total), d,
                                               C.byref(counter.c) if counter else None))
        self.n = int(n_total)

    def build_tree(self, counter=None):
        self._ck(self._L.nb_build_tree(self._ctx, C.byref(counter.c) if counter else None))

    def accel(self, counter=None):
        ax, ay = np.zeros(self.n, np.float32), np.zeros(self.n, np.float32)
        self._ck(self._L.nb_accel(self._ctx, ax.ctypes.data, ay.ctypes.data, C.byref(counter.c) if counter else None))
        return ax, ay

    def accel_direct(self, first=0, count=None):
        count = self.n - first if count is None else count
        ax, ay = np.zeros(count, np.float32), np.zeros(count, np.float32)
        sec = C.c_double(0.0)
        self._ck(self._L.nb_accel_direct(self._ctx, first, count, ax.ctypes.data, ay.ctypes.data, C.byref(sec)))
        return ax, ay, sec.value

    def measure_fp32(self, kind):
        r = C.c_double(0.0)
        self._ck(self._L.nb_measure_fp32(self._ctx, int(kind), C.byref(r)))
        return r.value

    def flush_l2(self):
        self._ck(self._L.nb_flush_l2(self._ctx))

    # ---- introspection ---------------------------------------------------------------------
    def tree_info(self):
        info = NbTreeInfo()
        self._ck(self._L.nb_get_tree_info(self._ctx, C.byref(info)))
        return info

    def tree(self):
        info = self.tree_info()
        arrays, view = _lib.alloc_tree_view(int(info.n_bodies), int(info.n_nodes))
        self._ck(self._L.nb_get_tree(self._ctx, C.byref(view)))
        arrays["info"] = info
        return arrays

    def generate_on_device(self, scene, seed, n):
        """Fill the body state on the device (any NB_SCENE_*): no host buffer, no upload."""
        self._ck(self._L.nb_scene_generate_device(self._ctx, int(scene), int(seed), int(n)))
        self.n = int(self._L.nb_body_count(self._ctx))

    def empty(self):
        """`QuadTree::empty()` (quad_tree.rs:66-89) as a statistic: the count it would return; changes nothing."""
        v = C.c_uint32()
        self._ck(self._L.nb_quad_empty(self._ctx, C.byref(v)))
        return v.value

    def prune(self):
        """`QuadTree::prune()` (quad_tree.rs:94-137) as a statistic: the count it would return; changes nothing."""
        v = C.c_uint32()
        self._ck(self._L.nb_quad_prune(self._ctx, C.byref(v)))
        return v.value

    def profile_step(self):
        """One step with an event behind every kernel: [(name, microseconds), ...] in launch order."""
        buf = (_lib.NbKernelTime * 64)()
        k = C.c_size_t()
        self._ck(self._L.nb_profile_step(self._ctx, buf, 64, C.byref(k)))
        return [(buf[i].name.decode(), float(buf[i].us)) for i in range(k.value)]

    def draw(self, render_h=RENDER_HEIGHT):
        """`draw(particles, frame)` (src/main.rs:41-72) into a headless RGBA8 frame."""
        frame = np.zeros((render_h, render_h, 4), np.uint8)
        self._ck(self._L.nb_draw(self._ctx, frame.ctypes.data, render_h))
        return frame

    def save(self, path, steps_done=0):
        self._ck(self._L.nb_save_snapshot(self._ctx, str(path).encode(), int(steps_done)))

    def load(self, path):
        sd = C.c_uint64(0)
        self._ck(self._L.nb_load_snapshot(self._ctx, str(path).encode(), C.byref(sd)))
        info_n = np.fromfile(str(path), dtype="<u8", count=3)[2]
        self.n = int(info_n)
        return sd.value

    # ---- multi-GPU ---------------------------------------------------------------------------
    @staticmethod
    def comm_unique_id():
        buf = np.zeros(_lib.NB_UNIQUE_ID_BYTES, np.uint8)
        rc = _lib.lib().nb_comm_unique_id(buf.ctypes.data)
        if rc != 0:
            raise NbError(rc, _lib.lib().nb_last_error(None).decode())
        return buf

    def comm_init(self, rank, n_ranks, uid):
        uid = np.ascontiguousarray(uid, dtype=np.uint8)
        self._ck(self._L.nb_comm_init(self._ctx, rank, n_ranks, uid.ctypes.data))

    def comm_peer_export(self):
        """This rank's block of CUDA IPC handles for the peer-store exchange (all-gather them in rank order)."""
        buf = np.zeros(_lib.NB_PEER_BLOCK_BYTES, np.uint8)
        self._ck(self._L.nb_comm_peer_export(self._ctx, buf.ctypes.data))
        return buf

    def comm_peer_import(self, blocks):
        """blocks: the n_ranks export blocks in rank order; switches the exchange from ncclAllGather to NVLink stores."""
        blocks = np.ascontiguousarray(blocks, dtype=np.uint8)
        self._ck(self._L.nb_comm_peer_import(self._ctx, blocks.ctypes.data))

    def comm_peer_enable(self, on):
        self._ck(self._L.nb_comm_peer_enable(self._ctx, 1 if on else 0))

    def owned_range(self):
        f, c = C.c_uint64(), C.c_uint64()
        self._ck(self._L.nb_comm_owned_range(self._ctx, C.byref(f), C.byref(c)))
        return f.value, c.value
