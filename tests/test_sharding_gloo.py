"""World-size-2 (and 3) CPU test of the multi-GPU step PROTOCOL over gloo: Morton-range sharding
(nb_shard_range, the library's own host rule), per-rank traversal + integrate of the owned slice in
tree order, exchange of the POSITIONS only (padded all-gather of x, y), velocities owned by the rank that owns the
slot (s // per) and fetched from there by whoever integrates the body, a final pull before the state is read, local
permutation of mass/id (round 2: integrate_push_kernel / pull_velocity_kernel).  The per-rank compute
is done by the oracle here (no GPU in this tier); the CUDA implementation of the same protocol is
checked bit for bit on real GPUs by tests/multi_gpu_check.py."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, n, steps, kind, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import nbody_simulation_b200 as nb
    import oracle_binding as ob

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    theta, dt = np.float32(0.6), np.float32(0.1)
    p = nb.generate_scene(nb.NB_SCENE_TWO_GALAXY, 9, n)
    # replicated slot-order state (slot order == tree order of the previous step)
    x, y, vx, vy = (p[k].copy() for k in ("x", "y", "vx", "vy"))
    m, ident = p["weight"].copy(), np.arange(n, dtype=np.uint32)
    per = -(-(-(-n // world)) // 32) * 32  # nb_shard_range: whole 32-body warps
    def gather(a, count):  # padded all-gather of the owned slices, as ncclAllGather / the peer pushes lay them out
        send = torch.zeros(per, dtype=torch.float32)
        send[:count] = torch.from_numpy(a)
        recv = [torch.zeros(per, dtype=torch.float32) for _ in range(world)]
        dist.all_gather(recv, send)
        return torch.cat(recv)[:n].numpy().copy()

    # velocities: after the upload every rank holds all of them; after a step only those of its own slots are valid
    # here (the others are poisoned on purpose: nothing may read them without fetching from the owner)
    first, count = nb.shard_range(n, rank, world)
    for _ in range(steps):
        T = ob.Tree(kind, x, y, m)
        order = T.export()["order"]  # tree position -> slot
        own = order[first:first + count]
        # "peer loads": the velocity of slot s lives on rank min(s // per, world - 1)
        vx_all, vy_all = gather(vx[first:first + count], count), gather(vy[first:first + count], count)
        ax, ay, _, _, _ = T.accel(float(theta), own, threads=1)
        nvx = np.float32(vx_all[own] + ax * dt)
        nvy = np.float32(vy_all[own] + ay * dt)
        nx = np.float32(x[own] + nvx * dt)
        ny = np.float32(y[own] + nvy * dt)
        x, y = gather(nx, count), gather(ny, count)          # positions travel
        vx = np.full(n, np.nan, np.float32)
        vy = np.full(n, np.nan, np.float32)
        vx[first:first + count], vy[first:first + count] = nvx, nvy   # velocities stay
        m, ident = m[order], ident[order]
    vx, vy = gather(vx[first:first + count], count), gather(vy[first:first + count], count)  # pull_velocity_kernel
    res = np.zeros((5, n), np.float32)
    res[0, ident], res[1, ident], res[2, ident], res[3, ident] = x, y, vx, vy
    res[4, ident] = m.astype(np.float32)
    np.save(os.path.join(out_dir, f"rank{rank}.npy"), res)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,kind_name", [(2, "quad"), (3, "lbvh")])
def test_sharded_step_equals_single_rank(tmp_path, world, kind_name):
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import nbody_simulation_b200 as nb
    import oracle_binding as ob

    kind = ob.KIND_PIPE_QUAD if kind_name == "quad" else ob.KIND_PIPE_LBVH
    n, steps = 4001, 3
    mp.spawn(_worker, args=(world, _free_port(), n, steps, kind, str(tmp_path)), nprocs=world, join=True)
    p = nb.generate_scene(nb.NB_SCENE_TWO_GALAXY, 9, n)
    w = ob.OracleWorld(p)
    for _ in range(steps):
        w.update(kind, 0.1, 0.6)
    x, y, vx, vy, m = w.get()
    ref = np.stack([x, y, vx, vy, m.astype(np.float32)])
    for r in range(world):
        got = np.load(os.path.join(str(tmp_path), f"rank{r}.npy"))
        assert np.array_equal(got.view(np.uint32), ref.view(np.uint32)), f"rank {r}"
