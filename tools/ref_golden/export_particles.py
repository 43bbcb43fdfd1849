#!/usr/bin/env python
"""Step 1 of tools/ref_golden: the particles of tests/golden/golden_small.npz as raw 24-byte records
(f32 x, y, vx, vy; u32 weight; u32 pad) for the Rust test to read, plus the parameters it should be built with."""
import json
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
G = np.load(os.path.join(HERE, "..", "..", "tests", "golden", "golden_small.npz"))
np.ascontiguousarray(G["particles"]).tofile(os.path.join(HERE, "particles.bin"))
with open(os.path.join(HERE, "particles.json"), "w") as f:
    json.dump({"n": int(len(G["particles"]) // 24), "theta": float(G["theta"]), "dt": float(G["dt"]),
               "steps": int(G["steps"]), "note": "set `const THETA: f32` in the reference's src/main.rs:35 to theta"}, f)
print("wrote particles.bin / particles.json")
