"""Second half of __graft_entry__.smoke(): the multi-body kernels vs the oracle —
  (a) a small 10-cube-stack batch (256 worlds: the fused single-launch kernel, step_world_kernel);
  (b) a 4 096-world 10-cube-stack batch: the STAGED pipeline the benchmark times (stage_prepare_kernel, the all-cube
      SAT kernel, stage_manifold_kernel, stage_solve_kernel), a sample of worlds compared with the oracle;
  (c) an 8-world C4 pile (mixed hulls + joints) forced onto the staged pipeline: stage_sat_kernel (mixed hulls).
Kept to a few frames each: the driver also runs this under ncu with a launch cap."""
import os

import numpy as np


def _staged_case(name, init, hulls, frames, sample, hull_ids=None, joints=None, force_staged=False):
    import constraint_solver_b200 as cs
    import oracle_lib as xo
    import oracle_world as xw
    from constraint_solver_b200 import scenes

    n, B = init.shape
    J = 0 if joints is None else joints.shape[1]
    cpu = xo.rigid38(init[sample]).astype(np.float32)
    xw.step_worlds(cpu, B, list(hulls), scenes.FRAME_DT, 20, frames, threads=xo.hardware_threads(),
                   hull_ids=None if hull_ids is None else hull_ids[sample], joints=None if joints is None else joints[sample])
    old = os.environ.get("XPBD_WORLD_PATH")
    if force_staged:
        os.environ["XPBD_WORLD_PATH"] = "staged"  # read by the library at every step call
    try:
        with cs.Worlds(n, B, device=0, body_body_contacts=True, joints_per_world=J) as w:
            for h in hulls:
                w.add_hull(h)
            if hull_ids is not None:
                w.set_hulls(hull_ids)
            if joints is not None:
                w.set_joints(joints)
            w.upload(init)
            w.integrate(scenes.FRAME_DT, 20, frames)
            st = w.stats()
            got = xo.rigid38(w.download()).reshape(n, B, 38)[sample].reshape(-1, 38)
    finally:
        if force_staged:
            if old is None:
                os.environ.pop("XPBD_WORLD_PATH", None)
            else:
                os.environ["XPBD_WORLD_PATH"] = old
    assert st["kernel_launches"] >= 4 * 20 * frames, "the staged pipeline did not run"
    assert st["overflow_worlds"] == 0 and st["nan_worlds"] == 0
    assert np.array_equal(got, cpu.astype(np.float64)), f"{name}: GPU state differs from the fp32 oracle"
    print(f"smoke ok ({name}, staged pipeline): {n} worlds x {frames} frames, {st['constraints']} constraints, "
          f"{st['kernel_launches']} launches, {st['elapsed_ms']:.3f} ms on device")


def smoke_world():
    import constraint_solver_b200 as cs
    import oracle_lib as xo
    import oracle_world as xw
    from constraint_solver_b200 import scenes

    cube = cs.Polytope.new_cube()
    n, frames = 256, 30
    init = scenes.c3_stack(n)
    cpu = xo.rigid38(init).astype(np.float32)
    c = xw.step_worlds(cpu, 10, [cube], scenes.FRAME_DT, 20, frames, threads=xo.hardware_threads())
    with cs.Worlds(n, 10, device=0, body_body_contacts=True) as w:
        w.add_hull(cube)
        w.upload(init)
        w.integrate(scenes.FRAME_DT, 20, frames)
        st = w.stats()
        got = xo.rigid38(w.download())
    assert st["overflow_worlds"] == 0
    assert np.array_equal(got, cpu.astype(np.float64)), "multi-body GPU state differs from the fp32 oracle"
    assert st["constraints"] == c["constraints"]
    print(f"smoke ok (10-cube stacks): {n} worlds x {frames} frames, {st['constraints']} constraints, {st['elapsed_ms']:.3f} ms on device")

    n = 4096
    _staged_case("C3 10-cube stacks", scenes.c3_stack(n), [cube], 3, np.arange(0, n, 64))
    init, hulls, ids, joints = scenes.c4_pile(8, spacing=1.5)  # dense lattice: candidate pairs from the first substep
    _staged_case("C4 mixed-hull piles + joints", init, hulls, 3, np.arange(8), hull_ids=ids, joints=joints, force_staged=True)
