#!/bin/bash
# compute-sanitizer on the smallest cases that exercise every kernel (one tool per gpurun call: B200_PROFILING.md)
set -u
TOOL=${1:-memcheck}
mkdir -p gpurun_out
python - > gpurun_out/sanitize_plain.log 2>&1 <<'PY'
import os, sys
sys.path.insert(0, "tests")
import numpy as np
import constraint_solver_b200 as cs
from constraint_solver_b200 import scenes
def run():
    cube = cs.Polytope.new_cube()
    with cs.Worlds(300) as w:                      # single-body kernel
        w.add_hull(cube); w.upload(scenes.c2_perturbed(300)); w.integrate(scenes.FRAME_DT, 25, 40)
        print("single", w.stats()["constraints"])
    for path in ("staged", "fused"):
        os.environ["XPBD_WORLD_PATH"] = path
        with cs.Worlds(37, 10, body_body_contacts=True) as w:   # 10-cube stacks, ragged world count
            w.add_hull(cube); w.upload(scenes.c3_stack(37)); w.integrate(scenes.FRAME_DT, 20, 4)
            print(path, "c3", w.stats()["constraints"], w.stats()["overflow_worlds"])
        init, hulls, ids, joints = scenes.c4_pile(5)
        with cs.Worlds(5, 64, body_body_contacts=True, joints_per_world=32) as w:   # mixed hulls + joints
            for h in hulls: w.add_hull(h)
            w.set_hulls(ids); w.set_joints(joints); w.upload(init); w.integrate(scenes.FRAME_DT, 20, 3)
            print(path, "c4", w.stats()["constraints"], w.stats()["overflow_worlds"])
run()
PY
echo "plain exit $?"; tail -5 gpurun_out/sanitize_plain.log
cat > /tmp/san.py <<'PY'
exec(open("gpurun_out/sanitize_driver.py").read())
PY
sed -n '/^import os/,/^run()/p' "$0" > gpurun_out/sanitize_driver.py
timeout 1500 compute-sanitizer --tool $TOOL --error-exitcode 3 python /tmp/san.py > gpurun_out/sanitize_$TOOL.log 2>&1
echo "$TOOL exit $?"; tail -12 gpurun_out/sanitize_$TOOL.log
