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
    ap.add_argument("--steps", type=int, default=20)
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
    ap.add_argument("--full-sort", action="store_true",
                    help="quadtree: sort every key digit in every build (default: only the digits the tree depth needs, "
                         "checked on the device each step; same tree either way)")
    return ap.parse_args()


def make_scene(nb, args, n_total):
    kinds = {"disc": nb.NB_SCENE_UNIFORM_DISC, "plummer": nb.NB_SCENE_PLUMMER, "galaxy": nb.NB_SCENE_TWO_GALAXY,
             "default": nb.NB_SCENE_DEFAULT}
    return nb.generate_scene(kinds[args.scene], 1, n_total)


def workload_name(args, n_total):
    return (f"{n_total}-body {args.scene} scene, {args.tree} tree, theta={args.theta}, dt=0.1 "
            f"(BASELINE.json configs[1] scaled weakly: {args.n_per_gpu} bodies per GPU)")


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


def cpu_port_steps(parts, args, n_steps, warmup, stride, threads=0):
    """The reference's CPU path (oracle port of quad_tree.rs / bvh_tree.rs + main.rs:388-425) on the host
    cores.  stride > 1: the traversal covers every stride-th body (rayon's per-body independence makes
    the sample exact for throughput); build and integrate are always full."""
    import oracle_binding as ob
    from nbody_simulation_b200._lib import NbCounting

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
    if rank != 0:
        return
    import nbody_simulation_b200 as nb

    n_total = args.n_per_gpu * world
    parts = make_scene(nb, args, n_total)
    n_total = len(parts)
    stride = max(1, world)
    r = cpu_port_steps(parts, args, args.steps, min(args.warmup, 1), stride)
    sample = (f"{args.steps} World::update steps of the oracle port (literal "
              f"{'quad_tree.rs + carried-over traversal' if args.tree == 'quad' else 'bvh_tree.rs'}) on {n_total} "
              f"bodies; build single-threaded, traversal OpenMP dynamic,5000 over every {stride}-th body"
              + (" (scaled by the stride)" if stride > 1 else ""))
    line = {
        "impl": "reference", "metric": "barnes_hut_body_interactions_per_s", "value": r["value"],
        "unit": "interactions/s", "n_gpus": world, "steps": args.steps, "warmup": min(args.warmup, 1),
        "ms_per_step": r["seconds_per_step"] * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(args, n_total), "n_bodies": n_total, "theta": args.theta,
                   "tree": args.tree},
        "steps_per_s": 1.0 / r["seconds_per_step"] if r["seconds_per_step"] > 0 else 0.0,
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

    n_total = args.n_per_gpu * world
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
        host, owner = nb.pinned_particles(n_total)
        host[:] = parts
        for _ in range(2):
            w.update_host(host, 0.1)
        host[:] = parts
        ce = nb.Counting()
        barrier()
        t1 = time.perf_counter()
        for _ in range(args.steps):
            w.update_host(host, 0.1, ce)
        barrier()
        e2e_s = allreduce(time.perf_counter() - t1, dist.ReduceOp.MAX if dist else None)
        w.set_counting(True)
        w.set_particles(parts)
        cr = nb.Counting()
        w.step(args.steps, cr)  # the e2e loop starts from `parts` (no warm-up steps): its own replay
        e2e_inter = allreduce(float(cr.interactions), dist.ReduceOp.SUM if dist else None)
        e2e = {"value": e2e_inter / e2e_s, "unit": "interactions/s", "h2d_bytes_per_step": 24 * n_total,
               "d2h_bytes_per_step": 24 * n_total, "ms_per_step": e2e_s / args.steps * 1e3,
               "steps_per_s": args.steps / e2e_s, "call": "nb_world_update (pinned host nb_particle[])"}
        del host, owner

    clocks = sampler.stop() if sampler else None
    # what the builds of the timed steps sorted (one more build of the final state, outside every timed region)
    w.build_tree()
    ti = w.tree_info()
    sort_note = ("n/a" if args.tree == "bvhref" else "every key digit" if ti.sort_top_digits == 0 else
                 f"top {ti.sort_top_digits} of 6 key digits per build, the depth guess checked on the device in every step "
                 f"and redone in full when too shallow (last build redone: {bool(ti.sort_redone)}); same tree bit for bit "
                 "as the full sort (tests/test_gpu_parity.py::test_top_digit_sort_*), --full-sort turns it off")
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
            "ms_per_step": dev_s / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(args, n_total), "n_bodies": n_total, "theta": args.theta,
                       "tree": args.tree, "force_mode": args.force, "parallelism": f"morton-range x{world}" + (f", exchange={args.exchange}" if world > 1 else ""),
                       "l2": "not flushed" if args.no_flush else "256 MB scratch overwritten between timed steps",
                       "interaction_count": "from an identical replay of the timed steps with counters on",
                       "sort": sort_note, "tree_depth": int(ti.max_depth)},
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
        if world == 1 and not args.no_cpu_baseline:
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


if __name__ == "__main__":
    main()
