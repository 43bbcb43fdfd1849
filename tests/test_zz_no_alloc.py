"""nb_accel_direct_noalloc (include/nbody_b200.h): the allocation-free variant of the tiled direct sum, whose fp64
accumulators live in a file-scope `static __device__` array instead of a per-call cudaMalloc.  No counterpart in the
reference.  Checked through the C ABI against the existing entry point
nb_accel_direct, on two calls in a row (the array keeps the previous call's contents: the second call must not see
them).  In a file of its own, collected last."""
import ctypes as C

import numpy as np
import pytest

import nbody_simulation_b200 as nb
from nbody_simulation_b200 import _lib

pytestmark = pytest.mark.gpu


def normwise_err(ax, ay, rx, ry):
    """||da||_2 / ||a||_2, the tolerance norm of the other GPU tests (tests/oracle_binding.py)."""
    ax, ay, rx, ry = (np.asarray(v, np.float64) for v in (ax, ay, rx, ry))
    return float(np.sqrt(np.sum((ax - rx) ** 2 + (ay - ry) ** 2)) / np.sqrt(np.sum(rx ** 2 + ry ** 2)))


def device_bodies(p):
    import torch

    x = torch.from_numpy(np.ascontiguousarray(p["x"])).cuda()
    y = torch.from_numpy(np.ascontiguousarray(p["y"])).cuda()
    w = torch.from_numpy(np.ascontiguousarray(p["weight"]).view(np.int32)).cuda()  # u32 bits
    return x, y, w


def call_noalloc(x, y, w, first, count, mode):
    import torch

    L = _lib.lib()
    ax = torch.full((count,), float("nan"), dtype=torch.float32, device="cuda")
    ay = torch.full((count,), float("nan"), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    rc = L.nb_accel_direct_noalloc(x.data_ptr(), y.data_ptr(), w.data_ptr(), x.numel(), first, count, mode,
                                   ax.data_ptr(), ay.data_ptr(), None)
    assert rc == 0, L.nb_last_error(None).decode()
    torch.cuda.synchronize()
    return ax.cpu().numpy(), ay.cpu().numpy()


@pytest.mark.parametrize("mode", [nb.NB_FORCE_FAST, nb.NB_FORCE_EXACT])
def test_noalloc_direct_sum_matches_nb_accel_direct_on_two_calls_in_a_row(mode):
    p = nb.generate_scene(nb.NB_SCENE_TWO_GALAXY, 11, 40_001)
    ref = nb.World(len(p), theta=0.5, force_mode=mode)
    ref.set_particles(p)
    x, y, w = device_bodies(p)
    # second call: other targets, fewer of them, so stale sums of the first call would sit under it; third: a repeat
    for first, count in ((0, 5000), (12_345, 3000), (0, 5000), (len(p) - 7, 7)):
        rx, ry, _ = ref.accel_direct(first, count)
        ax, ay = call_noalloc(x, y, w, first, count, mode)
        assert np.isfinite(ax).all() and np.isfinite(ay).all()
        # same tiles, same fp32 terms; only the order of the fp64 atomic adds across source chunks may differ.
        # 2e-6 normwise is the bar of test_direct_sum_against_fp64 for this kernel
        assert normwise_err(ax, ay, rx, ry) <= 2e-6, (first, count)
        assert np.max(np.abs(ax - rx)) <= 1e-5 * np.max(np.abs(rx)) and np.max(np.abs(ay - ry)) <= 1e-5 * np.max(np.abs(ry))
    ref.close()


def test_noalloc_rejects_more_targets_than_the_static_array_holds():
    p = nb.generate_scene(nb.NB_SCENE_UNIFORM_DISC, 3, _lib.NB_NOALLOC_MAX_TARGETS + 100)
    x, y, w = device_bodies(p)
    import torch

    L = _lib.lib()
    out = torch.zeros(_lib.NB_NOALLOC_MAX_TARGETS + 100, dtype=torch.float32, device="cuda")
    rc = L.nb_accel_direct_noalloc(x.data_ptr(), y.data_ptr(), w.data_ptr(), x.numel(), 0,
                                   _lib.NB_NOALLOC_MAX_TARGETS + 1, nb.NB_FORCE_FAST, out.data_ptr(), out.data_ptr(), None)
    assert rc == _lib.NB_ERR_CAPACITY
    assert b"NB_NOALLOC_MAX_TARGETS" in L.nb_last_error(None)
    # the maximum itself is accepted, and a range past the end of the bodies is not
    ax, ay = call_noalloc(x, y, w, 100, _lib.NB_NOALLOC_MAX_TARGETS, nb.NB_FORCE_FAST)
    assert np.isfinite(ax).all() and np.isfinite(ay).all()
    rc = L.nb_accel_direct_noalloc(x.data_ptr(), y.data_ptr(), w.data_ptr(), x.numel(), x.numel() - 5, 10,
                                   nb.NB_FORCE_FAST, out.data_ptr(), out.data_ptr(), None)
    assert rc == _lib.NB_ERR_CAPACITY
