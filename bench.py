#!/usr/bin/env python
"""bench.py — Barnes-Hut step throughput on B200 (BASELINE.json metric) with roofline, CPU baseline
and the reference arm.

  python bench.py [--gpus N] [--steps K] [--warmup W]            # this repo's CUDA path
  python bench.py --impl reference [--gpus N] --steps K --warmup W   # reference CPU path (oracle port)
  torchrun --nproc-per-node N bench.py --gpus N ...                # one rank per GPU

Workload at N=1 (BASELINE.json configs[1]): 1M-body uniform disc, quadtree (cap 8), theta = 0.5.
At N GPUs the run is WEAK-scaled: N x 1M bodies in the same domain, Morton-range sharded, slices
all-gathered with NCCL each step.  A "step" is one World::update (build + traverse + integrate).
`value` = body-interactions/s of the whole job, bodies resident in HBM; `e2e` = the same through
nb_world_update with the particle array in pinned HOST memory (H2D + step + D2H per step).
"""
import argparse
import ctypes
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

FLOP_PER_INTERACTION = 14  # literal count of src/main.rs:235-253


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200,
                    help="timed steps (default 200: a 0.2 s timed region at 1M bodies, so that the clock samples mean something)")
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--n-per-gpu", type=int, default=1_000_000)
    ap.add_argument("--scene", default="disc", choices=["disc", "plummer", "galaxy", "default"])
    ap.add_argument("--tree", default="quad", choices=["quad", "lbvh", "bvhref"])
    ap.add_argument("--theta", type=float, default=0.5)
    ap.add_argument("--force", default="fast", choices=["fast", "exact"])
    ap.add_argument("--exchange", default="peer", choices=["peer", "nccl"],
                    help="multi-GPU slice exchange: NVLink stores from the integrate kernel, or ncclAllGather")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-flush", action="store_true", help="do not flush L2 between timed steps")
    ap.add_argument("--config", type=int, default=0, choices=[0, 1, 2, 3, 4, 5],
                    help="BASELINE.json configs[N-1] (BASELINE.md §2 table): sets scene / bodies / tree / theta; "
                         "0 = the default line (config 2 weak-scaled).  1: default scene (CPU path of the reference: use "
                         "--impl reference; ours runs its own tree for comparison); 3: 4M Plummer, LBVH; "
                         "4: 16M two-galaxy over --gpus ranks (strong-scaled); 5: 64M Plummer over --gpus ranks")
    ap.add_argument("--no-parity", action="store_true", help="N > 1: skip the bit-identity check against one GPU")
    ap.add_argument("--full-sort", action="store_true",
                    help="quadtree: sort every key digit in every build (default: only the digits the tree depth needs, "
                         "checked on the device each step; same tree either way)")
    args = ap.parse_args()
    args.total_bodies = None  # None: weak scaling, n_per_gpu x world
    if args.config == 1:
        args.scene, args.total_bodies = "default", 0
        if args.tree == "quad" and "--tree" not in sys.argv:
            args.tree = "bvhref"  # the reference's live tree (bvh_tree.rs)
        if "--theta" not in sys.argv:
            args.theta = 50.0     # THETA as shipped (main.rs:35); pass --theta 0.5 for the other half of config 1
    elif args.config == 3:
        args.scene, args.tree, args.total_bodies = "plummer", ("lbvh" if "--tree" not in sys.argv else args.tree), 4_000_000
    elif args.config == 4:
        args.scene, args.total_bodies = "galaxy", 16_000_000
    elif args.config == 5:
        args.scene, args.total_bodies = "plummer", 64_000_000
    return args


def total_bodies(args, world):
    return args.n_per_gpu * world if args.total_bodies is None else args.total_bodies


def make_scene(nb, args, n_total):
    kinds = {"disc": nb.NB_SCENE_UNIFORM_DISC, "plummer": nb.NB_SCENE_PLUMMER, "galaxy": nb.NB_SCENE_TWO_GALAXY,
             "default": nb.NB_SCENE_DEFAULT}
    return nb.generate_scene(kinds[args.scene], 1, n_total)


def workload_name(args, n_total):
    which = {0: f"BASELINE.json configs[1] scaled weakly: {args.n_per_gpu} bodies per GPU", 1: "BASELINE.json configs[0]",
             2: "BASELINE.json configs[1]", 3: "BASELINE.json configs[2]", 4: "BASELINE.json configs[3], strong-scaled",
             5: "BASELINE.json configs[4], strong-scaled"}[args.config]
    return f"{n_total}-body {args.scene} scene, {args.tree} tree, theta={args.theta}, dt=0.1 ({which})"


def workload_config(args, n_total):
    """The `config` object: names the workload only, identical in both arms (ours / --impl reference)."""
    return {"workload": workload_name(args, n_total), "n_bodies": n_total, "scene": args.scene, "theta": args.theta,
            "tree": args.tree, "dt": 0.1}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}",
                                       "--format=csv,noheader,nounits", "-lms", "10"], stdout=self.f,
                                      stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def wait_first_sample(self, timeout=5.0):
        """nvidia-smi needs a few hundred ms to start; the timed region of the default run is shorter than that."""
        if self.p is None:
            return
        t0 = time.perf_counter()
        while time.perf_counter() - t0 < timeout:
            try:
                if os.path.getsize(self.f.name) > 0:
                    return
            except OSError:
                return
            time.sleep(0.01)

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.f.read().splitlines():
            parts = [s.strip() for s in line.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
            except ValueError:
                continue
            for nm, v in zip(names, parts[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        os.unlink(self.f.name)
        if sm:
            out.update(sm_mhz=statistics.median(sm), sm_max_mhz=max(mx), reasons=sorted(reasons), samples=len(sm))
        return out


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0}, "fallback"


def hbm_rooflines(kernel_times, n, n_nodes, hbm_gbs, peak_src):
    """Achieved HBM GB/s of the step's memory-bound kernels: ALGORITHMIC bytes (SURVEY.md §8 d) over the
    device time nb_profile_step measured for each (one step, events behind every kernel, L2 flushed before)."""
    per_body = {  # bytes per body
        "morton_kernel": 20.0,          # x, y read (8) + key (8) + slot index (4) written
        "morton_repair_kernel": 28.0,   # ... + the key the slot's body had at the previous build (8), sort_repair.cuh
        "sort_repair_kernel": 20.0,     # key read (8); key + slot written (12)
        "sort_histogram_kernel": 8.0,   # one read of the keys, the digits of the passes that run
        "sort_onesweep_kernel": 24.0,   # (key, slot) read + written
        "tree_count_kernel": 13.0,      # key read (8) + leaf depth (1) + node-count prefix (4) written
        "integrate_resort_kernel": 60.0,  # order 4 + acc 8 + 24 gathered + 24 written
    }
    per_node = {"tree_emit_kernel": 84.0, "tree_link_kernel": 56.0}  # node records written / linked
    out, seen = [], {}
    for name, us in kernel_times:
        if name in per_body:
            b = per_body[name] * n
        elif name in per_node:
            b = per_node[name] * n_nodes
        elif name == "tree_finish_kernel":
            b = 40.0 * n + 100.0 * n_nodes  # leaf fix-up + gather per body; summaries read / written per node
        else:
            continue
        k = seen.get(name, 0)
        seen[name] = k + 1
        gbs = b / (us * 1e-6) / 1e9 if us > 0 else 0.0
        out.append({"kernel": name + (f" #{k + 1}" if name == "sort_onesweep_kernel" else ""), "bytes": b, "us": us,
                    "achieved_gbs": gbs, "frac": gbs / hbm_gbs})
    return {"peak_gbs": hbm_gbs, "peak_source": f"MEASURED_PEAKS.json hbm_gbs ({peak_src})", "kernels": out,
            "all_kernels_us": [[nm, round(us, 2)] for nm, us in kernel_times]}


def cpu_port_steps(parts, args, n_steps, warmup, stride, threads=0):
    """The reference's CPU path (oracle port of quad_tree.rs / bvh_tree.rs + main.rs:388-425) on the host
    cores.  stride > 1: the traversal covers every stride-th body (rayon's per-body independence makes
    the sample exact for throughput); build and integrate are always full."""
    import oracle_binding as ob
    from nbody_simulation_b200._lib import NbCounting  # struct layout only; does not load libnbody_b200.so

    kind = ob.KIND_LITERAL_QUAD if args.tree == "quad" else ob.KIND_LITERAL_BVH
    if threads <= 0:  # all usable host cores, whatever OMP_NUM_THREADS says (torchrun exports OMP_NUM_THREADS=1)
        threads = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    ow = ob.OracleWorld(parts)
    for _ in range(warmup):
        ow.update(kind, 0.1, args.theta, sample_stride=stride, threads=threads)
    c = NbCounting()
    t0 = time.perf_counter()
    for _ in range(n_steps):
        ow.update(kind, 0.1, args.theta, sample_stride=stride, threads=threads, counting=c)
    wall = time.perf_counter() - t0
    est_s = c.build_s + c.traverse_s * stride + c.integrate_s
    est_inter = c.interactions * stride
    return {
        "value": est_inter / est_s if est_s > 0 else 0.0,
        "seconds_per_step": est_s / max(1, n_steps),
        "interactions_per_step": est_inter / max(1, n_steps),
        "build_s": c.build_s / max(1, n_steps), "traverse_s": c.traverse_s * stride / max(1, n_steps),
        "integrate_s": c.integrate_s / max(1, n_steps), "wall_s": wall,
        "cores": threads,
    }


def run_reference(args, rank, world):
    """The reference's CPU implementation of the path (oracle port; the reference is Rust and cannot be built here)
    on the host cores of this box: every body walked, every step timed; nothing of the product library is loaded."""
    if rank != 0:
        return
    os.environ.setdefault("NB_ORACLE_NATIVE", "1")  # build the oracle for THIS box's cores
    import oracle_binding as ob

    kinds = {"disc": 1, "plummer": 2, "galaxy": 3, "default": 0}  # NB_SCENE_* of include/nbody_b200.h
    n_total = total_bodies(args, world)
    parts = ob.generate_scene(kinds[args.scene], 1, n_total)
    n_total = len(parts)
    warm = max(3, args.warmup)
    # bound the run (~4 min): time one step, then run as many of the requested steps as fit; print what was run
    t0 = time.perf_counter()
    probe = cpu_port_steps(parts, args, 1, 0, 1)
    t_step = max(time.perf_counter() - t0, 1e-3)
    budget = 240.0
    warm_run = min(warm, max(0, int(0.2 * budget / t_step)))
    steps_run = max(1, min(args.steps, int((budget - (1 + warm_run) * t_step) / t_step)))
    r = cpu_port_steps(parts, args, steps_run, warm_run, 1)
    sample = (f"{steps_run} full World::update steps (every body walked) of the oracle port (literal "
              f"{'quad_tree.rs + carried-over traversal' if args.tree == 'quad' else 'bvh_tree.rs'}) on {n_total} "
              f"bodies; build single-threaded, traversal OpenMP dynamic,5000 on {r['cores']} threads; "
              f"g++ {ob.build_flags}")
    line = {
        "impl": "reference", "metric": "barnes_hut_body_interactions_per_s", "value": r["value"],
        "unit": "interactions/s", "n_gpus": world, "steps": steps_run, "steps_requested": args.steps, "warmup": warm,
        "warmup_run": warm_run,
        "ms_per_step": r["seconds_per_step"] * 1e3, "higher_is_better": True, "scaling": "weak" if args.total_bodies is None else "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args, n_total),
        "steps_per_s": 1.0 / r["seconds_per_step"] if r["seconds_per_step"] > 0 else 0.0,
        "interactions_per_step": r["interactions_per_step"],
        "cpu_baseline": {"value": r["value"], "unit": "interactions/s", "cores": r["cores"], "kind": "port",
                         "sample": sample},
        "e2e": {"value": r["value"], "unit": "interactions/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "phases_ms": {"build": r["build_s"] * 1e3, "traverse": r["traverse_s"] * 1e3,
                      "integrate": r["integrate_s"] * 1e3},
    }
    print(json.dumps(line), flush=True)


def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import numpy as np
    import torch

    import nbody_simulation_b200 as nb

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the CUDA path has no CPU fallback (use --impl reference)")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if dist is not None:
            dist.barrier(device_ids=[local_rank])
        torch.cuda.synchronize()

    def allreduce(v, op):
        if dist is None:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=op)
        return float(t.item())

    n_total = total_bodies(args, world)
    parts = make_scene(nb, args, n_total)
    n_total = len(parts)
    tree = {"quad": nb.NB_TREE_QUAD, "lbvh": nb.NB_TREE_LBVH, "bvhref": nb.NB_TREE_BVH_REF}[args.tree]
    force = nb.NB_FORCE_FAST if args.force == "fast" else nb.NB_FORCE_EXACT
    w = nb.World(n_total, theta=args.theta, dt=0.1, tree=tree, force_mode=force, device=local_rank,
                 full_sort=args.full_sort)
    if world > 1:
        uid = torch.zeros(nb._lib.NB_UNIQUE_ID_BYTES, dtype=torch.uint8, device="cuda")
        if rank == 0:
            uid.copy_(torch.from_numpy(nb.World.comm_unique_id()))
        dist.broadcast(uid, 0)
        w.comm_init(rank, world, uid.cpu().numpy())
        if args.exchange == "peer" and world <= 8:
            # fused integrate + exchange: every rank maps the peers' body arrays (CUDA IPC) and the integrate kernel
            # stores its slice into all of them over NVLink; the handle blocks travel through torch.distributed
            ok = 1.0
            try:
                mine = torch.from_numpy(w.comm_peer_export()).cuda()
            except nb.NbError as e:
                print(f"[bench] rank {rank}: peer export failed ({e}); falling back to ncclAllGather", file=sys.stderr)
                mine, ok = torch.zeros(nb._lib.NB_PEER_BLOCK_BYTES, dtype=torch.uint8, device="cuda"), 0.0
            blocks = [torch.empty_like(mine) for _ in range(world)]
            dist.all_gather(blocks, mine)
            if ok:
                try:
                    w.comm_peer_import(torch.stack(blocks).cpu().numpy())
                except nb.NbError as e:
                    print(f"[bench] rank {rank}: peer import failed ({e}); falling back to ncclAllGather", file=sys.stderr)
                    ok = 0.0
            all_ok = torch.tensor([ok], device="cuda")
            dist.all_reduce(all_ok, op=dist.ReduceOp.MIN)  # one exchange for everybody
            if all_ok.item() != 1.0:
                if ok:
                    w.comm_peer_enable(False)
                args.exchange = "nccl"
    w.set_particles(parts)

    # ---- interaction count of the timed steps: an identical (deterministic) replay with the diagnostic
    # counters compiled in; the timed runs below execute the production kernels (counters compiled out,
    # the reference never counted either).
    # the clock sampler runs from here to the end of the e2e loop: counter replay, warm-up, the timed steps and the
    # e2e steps are the same workload back to back (the timed region alone can be shorter than one sample period)
    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.wait_first_sample()
    barrier()  # rank 0 may have waited for nvidia-smi: nobody starts stepping (and waiting on peers) before it is back
    warm = max(3, args.warmup)
    w.step(warm)
    cc = nb.Counting()
    w.step(args.steps, cc)
    replay_inter = float(cc.interactions)
    w.set_particles(parts)
    w.set_counting(False)

    # ---- device-resident throughput -----------------------------------------------------------------
    w.step(warm)
    barrier()
    c = nb.Counting()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        if not args.no_flush:
            w.flush_l2()
        if dist is not None:
            # every rank starts the step together: the ranks hand-shake at the END of a step, so a rank whose host was
            # quicker through the flush would count its wait for the slowest starter as device time of the step
            barrier()
        w.step(1, c)
    barrier()
    wall = time.perf_counter() - t0
    dev_s = allreduce(c.total, dist.ReduceOp.MAX if dist else None)
    inter = allreduce(replay_inter, dist.ReduceOp.SUM if dist else None)
    rank_inter = replay_inter
    value = inter / dev_s

    # ---- end to end through the host-buffer drop-in call ----------------------------------------------
    e2e = None
    if not args.no_e2e:
        # N = 1: nb_world_update with the whole Vec.  N > 1: the Vec is sharded over the ranks' hosts
        # (nb_world_update_shard): each rank uploads and downloads only ITS 1/N slice, the slices travel between the
        # GPUs over NVLink.  Either way every body crosses PCIe once per direction per step.
        sharded = world > 1 and args.exchange == "peer" and args.tree != "bvhref"
        if sharded:
            lo, hi = rank * n_total // world, (rank + 1) * n_total // world
        else:
            lo, hi = 0, n_total
        host, owner = nb.pinned_particles(hi - lo)

        def e2e_step(counter=None):
            if sharded:
                w.update_host_shard(host, lo, n_total, 0.1, counter)
            else:
                w.update_host(host, 0.1, counter)

        host[:] = parts[lo:hi]
        for _ in range(2):
            e2e_step()
        host[:] = parts[lo:hi]
        ce = nb.Counting()
        barrier()
        t1 = time.perf_counter()
        for _ in range(args.steps):
            e2e_step(ce)
        barrier()
        e2e_s = allreduce(time.perf_counter() - t1, dist.ReduceOp.MAX if dist else None)
        e2e_digest = None
        if sharded:  # the sharded call must leave what the resident steps leave: compared below (parity)
            import hashlib
            e2e_digest = hashlib.sha256(host.tobytes()).hexdigest()
            e2e_final = host.copy()
        w.set_counting(True)
        w.set_particles(parts)
        cr = nb.Counting()
        w.step(args.steps, cr)  # the e2e loop starts from `parts` (no warm-up steps): its own replay
        e2e_inter = allreduce(float(cr.interactions), dist.ReduceOp.SUM if dist else None)
        per_rank = 24 * (hi - lo)
        e2e = {"value": e2e_inter / e2e_s, "unit": "interactions/s",
               "h2d_bytes_per_step": 24 * n_total if sharded or world == 1 else 24 * n_total * world,
               "d2h_bytes_per_step": 24 * n_total if sharded or world == 1 else 24 * n_total * world,
               "ms_per_step": e2e_s / args.steps * 1e3,
               "steps_per_s": args.steps / e2e_s,
               "call": ("nb_world_update_shard (each rank: its 1/N slice of the Vec in pinned host memory, "
                        f"{per_rank} B each way per step)" if sharded else "nb_world_update (pinned host nb_particle[])")}
        if sharded:
            mine_ok = bool(np.array_equal(w.particles[lo:hi].view(np.uint8), e2e_final.view(np.uint8)))
            e2e["same_state_as_resident_steps"] = bool(allreduce(1.0 if mine_ok else 0.0, dist.ReduceOp.MIN) == 1.0)
        del host, owner

    clocks = sampler.stop() if sampler else None

    # ---- N > 1: is the sharded run the SAME computation?  Every rank steps K times from `parts`, rank 0 repeats
    # the K steps on one GPU alone, and the SHA-256 digests of the full state (original body order) must agree.
    parity = None
    if world > 1 and not args.no_parity:
        import hashlib

        def digest(wd):
            x, y, vx, vy, m = wd.get_soa()
            return hashlib.sha256(x.tobytes() + y.tobytes() + vx.tobytes() + vy.tobytes() + m.tobytes()).digest()

        w.set_counting(False)
        w.set_particles(parts)
        barrier()
        w.step(args.steps)
        barrier()
        mine = torch.frombuffer(bytearray(digest(w)), dtype=torch.uint8).cuda()
        ref = torch.zeros(32, dtype=torch.uint8, device="cuda")
        if rank == 0:
            w1 = nb.World(n_total, theta=args.theta, dt=0.1, tree=tree, force_mode=force, device=local_rank,
                          full_sort=args.full_sort)
            w1.set_counting(False)
            w1.set_particles(parts)
            w1.step(args.steps)
            ref.copy_(torch.frombuffer(bytearray(digest(w1)), dtype=torch.uint8))
            w1.close()
        dist.broadcast(ref, 0)
        same = 1.0 if bool((mine == ref).all().item()) else 0.0
        same = allreduce(same, dist.ReduceOp.MIN)
        parity = {"bit_identical_to_1gpu": bool(same == 1.0), "steps": args.steps,
                  "what": "SHA-256 of x, y, vx, vy, weight of every body (original order) on every rank against a "
                          "single-GPU run of the same steps by rank 0"}

    # ---- per-kernel device times of one step (events behind every kernel), for the HBM-side rooflines
    kernel_times = None
    if world == 1:
        w.set_counting(False)
        w.set_particles(parts)
        w.step(warm)
        w.step(1)  # the depth is known: top-digit sort, movers counted
        w.step(1)  # ... and the builds the timed steps run (order repair when few bodies left their leaf cell)
        w.flush_l2()
        kernel_times = w.profile_step()
    # what the builds of the timed steps sorted (one more build of the final state, outside every timed region)
    w.build_tree()
    ti = w.tree_info()
    sort_note = ("n/a" if args.tree == "bvhref" else "every key digit" if ti.sort_top_digits == 0 else
                 f"{ti.sort_top_digits} passes over the top {ti.sort_top_levels} quadtree levels of the keys per build (a full sort takes 6), the depth guess checked on the device in every step "
                 f"and redone in full when too shallow (last build redone: {bool(ti.sort_redone)}); same tree bit for bit "
                 "as the full sort (tests/test_gpu_parity.py::test_top_digit_sort_*), --full-sort turns it off; a build whose "
                 "state is still in the previous tree's order re-inserts the bodies that left their leaf cell instead of "
                 f"sorting while those are few (csrc/sort_repair.cuh; last build: repaired={bool(ti.sort_repaired)}, "
                 f"{int(ti.sort_movers)} movers; NB_B200_NO_REPAIR=1 turns it off)")
    if clocks is not None:
        clocks["window"] = "counter replay, warm-up, timed steps and e2e steps (nvidia-smi -lms 10, median)"
    if rank == 0:
        peaks, peak_src = measured_peaks()
        fmax = float(peaks.get("sm_max_mhz", 1965.0))
        fp32_nominal = 148 * 128 * 2 * fmax * 1e6 / 1e12  # TFLOP/s
        fp32_peak = w.measure_fp32(0) / 1e12              # measured now: scalar FFMA chains (burst)
        pair_ceiling = w.measure_fp32(2)                  # pair force on register operands, interactions/s
        trav_s = c.sum_gravity / args.steps
        trav_tflops = FLOP_PER_INTERACTION * (rank_inter / args.steps) / trav_s / 1e12
        traffic = None
        try:
            with open(os.path.join(ROOT, "profiles", "traverse_traffic.json")) as f:
                traffic = json.load(f).get("dram_bytes_per_launch")
        except Exception:
            pass
        line = {
            "metric": "barnes_hut_body_interactions_per_s", "value": value, "unit": "interactions/s",
            "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup),
            "ms_per_step": dev_s / args.steps * 1e3, "higher_is_better": True,
            "scaling": "weak" if args.total_bodies is None else "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, n_total),
            "details": {"force_mode": args.force, "parallelism": f"morton-range x{world}" + (f", exchange={args.exchange}" if world > 1 else ""),
                        "l2": ("not flushed" if args.no_flush else "256 MB scratch overwritten between timed steps") +
                              (" (N > 1: then a barrier, so that every rank starts the step together)" if world > 1 else ""),
                        "interaction_count": "from an identical replay of the timed steps with counters on",
                        "sort": sort_note, "tree_depth": int(ti.max_depth), "n_nodes": int(ti.n_nodes)},
            "steps_per_s": args.steps / dev_s,
            "interactions_per_step": inter / args.steps,
            "phases_ms": {"build": c.build_bvh / args.steps * 1e3, "traverse": trav_s * 1e3,
                          "integrate": c.post_calculations / args.steps * 1e3,
                          "exchange": c.exchange / args.steps * 1e3},
            "wall_ms_per_step_incl_flush": wall / args.steps * 1e3,
            "clocks": clocks,
            "gpu_launches": int(c.launches),
            "roofline": {"bound": "fp32", "kernel": "traverse_fast_kernel" if args.force == "fast" else "traverse_kernel",
                         "achieved": trav_tflops,
                         "peak": fp32_peak, "unit": "TFLOP/s", "frac": trav_tflops / fp32_peak, "traffic": traffic,
                         "peak_source": ("FFMA microbenchmark (nb_measure_fp32 kind 0) measured in this run; "
                                         f"nominal 148 SM x 128 lanes x 2 flop x {fmax:.0f} MHz = {fp32_nominal:.1f}; "
                                         "MEASURED_PEAKS.json holds no fp32 entry"),
                         "flop_per_interaction": FLOP_PER_INTERACTION,
                         "pair_force_ceiling_interactions_per_s": pair_ceiling,
                         "frac_of_pair_force_ceiling": (rank_inter / args.steps) / trav_s / pair_ceiling},
        }
        if e2e:
            line["e2e"] = e2e
        if parity is not None:
            line["parity"] = parity
        if kernel_times:
            line["roofline_hbm"] = hbm_rooflines(kernel_times, n_total, int(ti.n_nodes), float(peaks.get("hbm_gbs", 6650.0)),
                                                 peak_src)
        if world == 1 and not args.no_cpu_baseline:
            os.environ.setdefault("NB_ORACLE_NATIVE", "1")  # oracle built for this box's cores
            r = cpu_port_steps(parts, args, 3, 1, 1)
            line["cpu_baseline"] = {
                "value": r["value"], "unit": "interactions/s", "cores": r["cores"], "kind": "port",
                "sample": (f"3 full World::update steps of the oracle port (literal "
                           f"{'quad_tree.rs' if args.tree == 'quad' else 'bvh_tree.rs'}) on the same {n_total} bodies; "
                           f"{r['seconds_per_step']:.2f} s/step: build {r['build_s']:.2f} s single-threaded, "
                           f"traversal {r['traverse_s']:.2f} s on {r['cores']} threads"),
                "steps_per_s": 1.0 / r["seconds_per_step"]}
        print(json.dumps(line), flush=True)
    w.close()
    if dist is not None:
        dist.destroy_process_group()
    if parity is not None and not parity["bit_identical_to_1gpu"]:
        raise SystemExit(3)


if __name__ == "__main__":
    main()
