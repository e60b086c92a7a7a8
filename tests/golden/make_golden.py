"""Generates tests/golden/c1_{f32,f64}.json from the CPU oracle (hex floats, bit-exact).

The reference (Rust) cannot be built or imported in this environment and ships no golden vectors, so these
fixtures are regression anchors for the oracle itself (PARITY UNPINNED)."""
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import oracle_lib as xo  # noqa: E402

FRAMES = [1, 10, 30, 60, 100, 300, 600]
for name, dtype in (("c1_f32", np.float32), ("c1_f64", np.float64)):
    w = xo.world_new().astype(dtype)
    cps, frame = [], 0
    for f in FRAMES:
        xo.world_integrate(w, 1 / 60, f - frame)
        frame = f
        cps.append({"frame": f, "a": [float(x).hex() for x in w[0]], "b": [float(x).hex() for x in w[1]]})
    with open(os.path.join(HERE, name + ".json"), "w") as fh:
        json.dump({"scene": "World::new (src/world.rs:12-31), dt=1/60, 25 substeps, both bodies vs cube",
                   "layout": "rigid38", "checkpoints": cps}, fh, indent=0)
    print(name, "written")
