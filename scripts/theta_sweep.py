#!/usr/bin/env python
"""theta_sweep.py — BASELINE.json configs[4]: accuracy vs throughput over theta on a large Plummer scene.

For every theta: steps/s and body-interactions/s of the whole step (CUDA events inside nb_step, counters
compiled out), the phase split, the traversal's fraction of the measured FP32 peak (14 flop per interaction,
SURVEY.md §8 d), and the normwise force error ||a_tree - a_direct||_2 / ||a_direct||_2 on a body sample against
the tiled O(N^2) direct sum on the device (K6, itself within 2e-6 of the fp64 oracle: tests/test_gpu_parity.py).

The scene is generated on the device (nb_scene_generate_device); nothing here touches oracle/ or torch.
One JSON object per theta on stdout, a summary table on stderr.

  python scripts/theta_sweep.py                       # 64M-body Plummer, quadtree, theta 0.3 .. 0.8, one GPU
  python scripts/theta_sweep.py --n 4000000 --tree lbvh
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import nbody_simulation_b200 as nb  # noqa: E402

FLOP_PER_INTERACTION = 14


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=64_000_000)
    ap.add_argument("--scene", default="plummer", choices=["plummer", "disc"])
    ap.add_argument("--tree", default="quad", choices=["quad", "lbvh"])
    ap.add_argument("--thetas", default="0.3,0.4,0.5,0.6,0.7,0.8")
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=1)
    ap.add_argument("--sample", type=int, default=16384, help="bodies of the direct-sum error sample")
    ap.add_argument("--seed", type=int, default=1)
    args = ap.parse_args()

    scene = {"plummer": nb.NB_SCENE_PLUMMER, "disc": nb.NB_SCENE_UNIFORM_DISC}[args.scene]
    tree = {"quad": nb.NB_TREE_QUAD, "lbvh": nb.NB_TREE_LBVH}[args.tree]
    thetas = [float(t) for t in args.thetas.split(",")]
    n = args.n
    w = nb.World(n, theta=thetas[0], dt=0.1, tree=tree, force_mode=nb.NB_FORCE_FAST)
    fp32_peak = w.measure_fp32(0)  # flop/s, scalar FFMA chains, measured now

    # the direct-sum yardstick for the sample: bodies [0, sample) in generation order are independent draws
    w.generate_on_device(scene, args.seed, n)
    ns = min(args.sample, n)
    t0 = time.perf_counter()
    rx, ry, direct_s = w.accel_direct(0, ns)
    print(f"[sweep] direct sum: {ns} targets x {n} sources in {direct_s:.3f} s "
          f"({ns * n / max(direct_s, 1e-9):.3e} interactions/s; wall {time.perf_counter() - t0:.1f} s)", file=sys.stderr)
    ref_norm = float(np.sqrt(np.sum(rx.astype(np.float64) ** 2 + ry.astype(np.float64) ** 2)))

    rows = []
    for theta in thetas:
        w.set_params(theta=theta)
        # forces of the initial state (one counted traversal) -> error against the direct sum
        w.generate_on_device(scene, args.seed, n)
        w.set_counting(True)
        ax, ay = w.accel()
        d = np.sqrt(np.sum((ax[:ns].astype(np.float64) - rx) ** 2 + (ay[:ns].astype(np.float64) - ry) ** 2))
        err = float(d / ref_norm)
        del ax, ay
        info = w.tree_info()
        # interaction count of the timed steps: identical replay with the counters on
        w.step(args.warmup)
        cc = nb.Counting()
        w.step(args.steps, cc)
        inter = float(cc.interactions)
        tests = float(cc.node_tests)
        # timed: production kernels (counters compiled out)
        w.generate_on_device(scene, args.seed, n)
        w.set_counting(False)
        w.step(args.warmup)
        c = nb.Counting()
        w.step(args.steps, c)
        k = args.steps
        trav_s = c.sum_gravity / k
        row = {
            "workload": f"{n}-body {args.scene}, {args.tree} tree, dt=0.1, one B200", "theta": theta,
            "steps_per_s": k / c.total, "ms_per_step": c.total / k * 1e3,
            "interactions_per_step": inter / k, "interactions_per_s": inter / c.total,
            "interactions_per_body": inter / k / n, "node_tests_per_body": tests / k / n,
            "phases_ms": {"build": c.build_bvh / k * 1e3, "traverse": trav_s * 1e3,
                          "integrate": c.post_calculations / k * 1e3},
            "traverse_tflops": FLOP_PER_INTERACTION * inter / k / trav_s / 1e12,
            "traverse_frac_fp32_peak": FLOP_PER_INTERACTION * inter / k / trav_s / fp32_peak,
            "fp32_peak_tflops_measured": fp32_peak / 1e12,
            "force_err_vs_direct": err, "err_sample": ns,
            "n_nodes": int(info.n_nodes), "max_depth": int(info.max_depth),
            "gpu_launches_per_step": c.launches / k,
        }
        rows.append(row)
        print(json.dumps(row), flush=True)

    print(f"\n| theta | ms/step | steps/s | interactions/body | interactions/s | build ms | traverse ms | "
          f"traverse frac of FP32 peak | force error vs direct ({ns} bodies) |", file=sys.stderr)
    print("|---:|---:|---:|---:|---:|---:|---:|---:|---:|", file=sys.stderr)
    for r in rows:
        print(f"| {r['theta']} | {r['ms_per_step']:.1f} | {r['steps_per_s']:.2f} | {r['interactions_per_body']:.0f} | "
              f"{r['interactions_per_s']:.3e} | {r['phases_ms']['build']:.2f} | {r['phases_ms']['traverse']:.1f} | "
              f"{r['traverse_frac_fp32_peak']:.3f} | {r['force_err_vs_direct']:.2e} |", file=sys.stderr)
    w.close()


if __name__ == "__main__":
    main()
