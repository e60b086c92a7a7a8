"""Second half of __graft_entry__.smoke(): a small 10-cube-stack batch through the multi-body kernel vs the oracle."""
import numpy as np


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
