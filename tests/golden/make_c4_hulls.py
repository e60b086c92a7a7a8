"""Writes tests/golden/c4_hulls.json: the hull table of config C4 (constraint_solver_b200.scenes.c4_hulls(): the
reference's three built-in hulls + five QuickHull hulls of seeded points, computed by the product's host-side
xpbd_quickhull).  The CPU reference arm of bench.py (`--impl reference`) and the oracle-side scene generator
(tests/oracle_scenes.py) load this fixture instead of importing the product, so that the reference arm never maps
libxpbd_b200.so.  tests/test_oracle_scenes.py checks the fixture against a fresh c4_hulls().  Floats are stored as hex."""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from constraint_solver_b200 import scenes  # noqa: E402

out = []
for h in scenes.c4_hulls():
    out.append({"vertices": [[float(x).hex() for x in v] for v in h.vertices], "edges": h.edges.tolist(), "faces": h.faces,
                "centroid": [float(x).hex() for x in h.centroid]})
with open(os.path.join(HERE, "c4_hulls.json"), "w") as f:
    json.dump(out, f, indent=0)
print("wrote", len(out), "hulls")
