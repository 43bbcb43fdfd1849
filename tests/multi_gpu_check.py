"""Multi-GPU parity check (run under torchrun on a box with >= 2 GPUs; not collected by pytest):

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
      --master-port 29511 tests/multi_gpu_check.py

Every rank steps the same scene with Morton-range sharding + NCCL all-gather and must end with
state bit-identical to a single-GPU run (and therefore to the oracle, see test_gpu_parity.py)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import nbody_simulation_b200 as nb  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ok = True
    for tree, n, peer in [(t, n, p) for t in (nb.NB_TREE_QUAD, nb.NB_TREE_LBVH, nb.NB_TREE_BVH_REF)
                          for n in (100_003, 7) for p in (False, True)]:
        if True:
            p = nb.generate_scene(nb.NB_SCENE_TWO_GALAXY, 3, n)
            steps = 5
            w = nb.World(len(p), theta=0.6, dt=0.1, tree=tree, force_mode=nb.NB_FORCE_EXACT, device=local)
            uid = torch.zeros(nb._lib.NB_UNIQUE_ID_BYTES, dtype=torch.uint8, device="cuda")
            if rank == 0:
                uid.copy_(torch.from_numpy(nb.World.comm_unique_id()))
            dist.broadcast(uid, 0)
            w.comm_init(rank, world, uid.cpu().numpy())
            if peer:  # NVLink stores from the integrate kernel instead of ncclAllGather
                mine = torch.from_numpy(w.comm_peer_export()).cuda()
                blocks = [torch.empty_like(mine) for _ in range(world)]
                dist.all_gather(blocks, mine)
                w.comm_peer_import(torch.stack(blocks).cpu().numpy())
            w.set_particles(p)
            c = nb.Counting()
            w.step(steps, c)
            q = w.particles
            ref = nb.World(len(p), theta=0.6, dt=0.1, tree=tree, force_mode=nb.NB_FORCE_EXACT, device=local)
            ref.set_particles(p)
            rc = nb.Counting()
            ref.step(steps, rc)
            r = ref.particles
            same = all(np.array_equal(q[k].view(np.uint32), r[k].view(np.uint32)) for k in ("x", "y", "vx", "vy", "weight"))
            tot = torch.tensor([float(c.interactions)], dtype=torch.float64, device="cuda")
            dist.all_reduce(tot)
            counts_ok = int(tot.item()) == rc.interactions
            f, cnt = w.owned_range()
            print(f"rank {rank}/{world} tree={tree} n={len(p)} exchange={'peer' if peer else 'nccl'} owned=[{f},{f + cnt}) bit-identical={same} "
                  f"interactions sum={int(tot.item())} single={rc.interactions} ok={counts_ok} "
                  f"exchange_ms/step={c.exchange / steps * 1e3:.3f}", flush=True)
            ok = ok and same and counts_ok
            w.close()
            ref.close()
    # ---- round 2: the sharded drop-in call (each rank holds 1/G of the Vec on its host) and the group walk ----------
    for tree, n, mode, grp in [(nb.NB_TREE_QUAD, 100_003, nb.NB_FORCE_EXACT, False), (nb.NB_TREE_LBVH, 50_001, nb.NB_FORCE_FAST, False),
                               (nb.NB_TREE_QUAD, 100_003, nb.NB_FORCE_FAST, True), (nb.NB_TREE_QUAD, 5, nb.NB_FORCE_EXACT, False)]:
        p = nb.generate_scene(nb.NB_SCENE_TWO_GALAXY, 4, n)
        steps = 4
        w = nb.World(len(p), theta=0.6, dt=0.1, tree=tree, force_mode=mode, device=local, fast_walk_group=grp)
        uid = torch.zeros(nb._lib.NB_UNIQUE_ID_BYTES, dtype=torch.uint8, device="cuda")
        if rank == 0:
            uid.copy_(torch.from_numpy(nb.World.comm_unique_id()))
        dist.broadcast(uid, 0)
        w.comm_init(rank, world, uid.cpu().numpy())
        mine = torch.from_numpy(w.comm_peer_export()).cuda()
        blocks = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(blocks, mine)
        w.comm_peer_import(torch.stack(blocks).cpu().numpy())
        lo, hi = rank * len(p) // world, (rank + 1) * len(p) // world
        host = p[lo:hi].copy()
        for _ in range(steps):
            w.update_host_shard(host, lo, len(p), 0.1)
        ref = nb.World(len(p), theta=0.6, dt=0.1, tree=tree, force_mode=mode, device=local, fast_walk_group=grp)
        ref.set_particles(p)
        ref.step(steps)
        r = ref.particles[lo:hi]
        same = all(np.array_equal(host[k].view(np.uint32), r[k].view(np.uint32)) for k in ("x", "y", "vx", "vy", "weight"))
        # and the resident multi-GPU steps with the same walk
        w.set_particles(p)
        w.step(steps)
        q = w.particles[lo:hi]
        same2 = all(np.array_equal(q[k].view(np.uint32), r[k].view(np.uint32)) for k in ("x", "y", "vx", "vy", "weight"))
        print(f"rank {rank}/{world} sharded nb_world_update_shard tree={tree} n={len(p)} mode={mode} group_walk={grp} "
              f"slice=[{lo},{hi}) bit-identical to 1 GPU: host slices {same}, resident steps {same2}", flush=True)
        ok = ok and same and same2
        w.close()
        ref.close()
    # ---- the fused step (traverse_fast_push_kernel: walk + integrate + NVLink push in one kernel), including a depth
    # guess that fails inside a batch of steps (the kernel then carries the state over and the host redoes the steps)
    for tree in (nb.NB_TREE_QUAD, nb.NB_TREE_LBVH):
        shallow = nb.generate_scene(nb.NB_SCENE_UNIFORM_DISC, 5, 20_000)
        deep = nb.generate_scene(nb.NB_SCENE_PLUMMER, 6, 150_001)
        res = []
        for multi in (True, False):
            w = nb.World(len(deep), theta=0.5, dt=0.1, tree=tree, force_mode=nb.NB_FORCE_FAST, device=local)
            if multi:
                uid = torch.zeros(nb._lib.NB_UNIQUE_ID_BYTES, dtype=torch.uint8, device="cuda")
                if rank == 0:
                    uid.copy_(torch.from_numpy(nb.World.comm_unique_id()))
                dist.broadcast(uid, 0)
                w.comm_init(rank, world, uid.cpu().numpy())
                mine = torch.from_numpy(w.comm_peer_export()).cuda()
                blocks = [torch.empty_like(mine) for _ in range(world)]
                dist.all_gather(blocks, mine)
                w.comm_peer_import(torch.stack(blocks).cpu().numpy())
            w.set_particles(shallow)
            w.step(2)
            w.set_particles(deep)  # much deeper: the first build's guess fails, steps 1..6 of the batch are redone
            c = nb.Counting()
            w.step(6, c)
            res.append((w.particles, c.interactions))
            w.close()
        (q, ni), (r, ni1) = res
        tot = torch.tensor([float(ni)], dtype=torch.float64, device="cuda")
        dist.all_reduce(tot)
        same = all(np.array_equal(q[k].view(np.uint32), r[k].view(np.uint32)) for k in ("x", "y", "vx", "vy", "weight"))
        print(f"rank {rank}/{world} fused walk + push tree={tree}: bit-identical to 1 GPU {same}, interactions {int(tot.item())} == {ni1}: "
              f"{int(tot.item()) == ni1}", flush=True)
        ok = ok and same and int(tot.item()) == ni1
    flag = torch.tensor([1.0 if ok else 0.0], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    dist.destroy_process_group()
    if flag.item() != 1.0:
        raise SystemExit("MULTI-GPU PARITY FAILED")
    if rank == 0:
        print("multi-gpu parity ok")


if __name__ == "__main__":
    main()
