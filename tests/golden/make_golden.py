"""Generates tests/golden/golden_small.npz from the CPU oracle (the reference is Rust and cannot run here,
so these are ORACLE-generated vectors: they pin the oracle and the CUDA path against regressions,
not against the reference binary — see DESIGN.md "parity unpinned").

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))
import nbody_simulation_b200 as nb  # noqa: E402
import oracle_binding as ob  # noqa: E402

N, THETA, DT, STEPS = 3000, 0.5, 0.1, 3


def scene():
    p = nb.generate_scene(nb.NB_SCENE_DEFAULT, 42)
    sel = np.r_[0:2, 2:len(p):len(p) // (N - 2)][:N]
    return np.ascontiguousarray(p[sel])


def build():
    p = scene()
    out = {"particles": p.view(np.uint8), "theta": np.float32(THETA), "dt": np.float32(DT), "steps": np.int32(STEPS)}
    for name, kind in (("quad", ob.KIND_PIPE_QUAD), ("lbvh", ob.KIND_PIPE_LBVH), ("bvhref", ob.KIND_LITERAL_BVH)):
        T = ob.Tree(kind, p["x"], p["y"], p["weight"])
        t = T.export()
        for k, v in t.items():
            if k != "info":
                out[f"{name}_{k}"] = v
        ax, ay, ni, nt, _ = T.accel(THETA, literal=(kind == ob.KIND_LITERAL_BVH))  # bvhref: recursive main.rs:349-386
        out[f"{name}_ax"], out[f"{name}_ay"] = ax, ay
        out[f"{name}_interactions"], out[f"{name}_tests"] = np.uint64(ni), np.uint64(nt)
        w = ob.OracleWorld(p)
        for _ in range(STEPS):
            w.update(kind, DT, THETA)
        x, y, vx, vy, _ = w.get()
        out[f"{name}_traj"] = np.stack([x, y, vx, vy])
    return out


if __name__ == "__main__":
    np.savez_compressed(os.path.join(HERE, "golden_small.npz"), **build())
    print("wrote", os.path.join(HERE, "golden_small.npz"))
