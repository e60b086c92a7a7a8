"""Single-call drop-ins (xpbd_solver_step, xpbd_world_integrate) and the headless driver with its rewind buffer
 against the fp32 oracle."""
import numpy as np
import pytest

import oracle_lib as xo
import constraint_solver_b200 as cs
from constraint_solver_b200 import scenes, solver

pytestmark = pytest.mark.gpu
DT = scenes.FRAME_DT


def _oracle_body(rec, hull_vertices, frames, substeps=25):
    r = xo.rigid38(np.array([rec])).astype(np.float32)
    xo.step_bodies(r, hull_vertices, DT, substeps, frames, threads=1)
    return r.astype(np.float64)


DYN = np.r_[22:28, 31:38]   # velocity, angular_velocity, position, rotation
CONST = np.r_[0:22, 28:31]  # inverse mass / inertia, forces, torques, centre of mass


def _same_dynamic_state_constants_untouched(rec, rec0, oracle38):
    """The drop-ins write back what solver::step mutates (fp32 values, exact in f64) and leave the caller's f64
    constants as they were (they used to come back fp32-rounded)."""
    got, orig = xo.rigid38(np.array([rec]))[0], xo.rigid38(np.array([rec0]))[0]
    return np.array_equal(got[DYN], oracle38.reshape(-1)[DYN]) and np.array_equal(got[CONST], orig[CONST])


def test_solver_step_matches_oracle_frame_by_frame():
    """solver::step(&mut rigid, &polytope, dt, substep_count) (src/solver.rs:3), called the way app.rs does."""
    cube = cs.Polytope.new_cube()
    w = scenes.default_world()
    rigid = w.a.copy()
    rec0 = rigid.rec.copy()
    ref = xo.rigid38(np.array([rigid.rec])).astype(np.float32)
    for _ in range(40):
        solver.step(rigid, cube, DT, 25)
        xo.step_bodies(ref, cube.vertices, DT, 25, 1, threads=1)
        # the one-shot call round-trips through f64 host records every frame; fp32 values survive that exactly
        assert _same_dynamic_state_constants_untouched(rigid.rec, rec0, ref.astype(np.float64))


def test_world_integrate_steps_both_bodies_against_p1():
    """World::integrate (src/world.rs:34-43): b is stepped against the CUBE hull (p1), not its own tetrahedron."""
    p1, p2 = cs.Polytope.new_cube(), 0.5 * cs.Polytope.new_tetrahedron()
    w = cs.World(p1, p2)
    a0, b0 = w.a.rec.copy(), w.b.rec.copy()
    frames = 100  # b is thrown upwards and lands after ~64 frames; only then does the hull matter
    for _ in range(frames):
        w.integrate(DT, p1, p2, None)
    assert _same_dynamic_state_constants_untouched(w.a.rec, a0, _oracle_body(a0, p1.vertices, frames))
    assert _same_dynamic_state_constants_untouched(w.b.rec, b0, _oracle_body(b0, p1.vertices, frames))
    assert not _same_dynamic_state_constants_untouched(w.b.rec, b0, _oracle_body(b0, p2.vertices, frames))
    # b's inverse inertia (537.6 / -76.8, src/world.rs:27) is not fp32-exact: it must still be the caller's f64 value
    assert np.array_equal(w.b.rec["inverse_inertia"], b0["inverse_inertia"]) and \
        not np.array_equal(b0["inverse_inertia"], b0["inverse_inertia"].astype(np.float32).astype(np.float64))


def test_one_shot_calls_validate_arguments():
    cube = cs.Polytope.new_cube()
    r = scenes.default_world().a
    with pytest.raises(cs.XpbdError):
        solver.step(r, cube, DT, 0)
    with pytest.raises(cs.XpbdError):
        solver.step(r, cube, -1.0, 25)


def test_headless_driver_history_and_rewind(tmp_path):
    """app.rs:204-215 without the window: every frame pushed; Down-arrow rewinds (app.rs:146-148)."""
    n = 64
    init = scenes.c2_perturbed(n)
    cube = cs.Polytope.new_cube()
    with cs.Worlds(n) as w:
        w.add_hull(cube)
        w.upload(init)
        d = cs.HeadlessDriver(w, DT, 25)
        d.step(12)
        assert len(d.states) == 13 and d.current_state == 12
        after12 = d.states[12].copy()
        d.rewind(5)
        assert d.current_state == 7
        assert np.array_equal(w.download_dynamic(), d.states[7])
        d.step(5)  # replaying from frame 7 reproduces frame 12 bit for bit
        assert np.array_equal(d.states[12], after12)
        traj = d.trajectory(world=3)
        assert traj.shape == (13, 13)
        # frame k of the history equals the oracle after k frames
        ref = xo.rigid38(init[3:4]).astype(np.float32)
        xo.step_bodies(ref, cube.vertices, DT, 25, 12, threads=1)
        assert np.array_equal(traj[12][:3], ref[0, 31:34])
        d.dump_csv(str(tmp_path / "t.csv"), world=3)
        rows = (tmp_path / "t.csv").read_text().strip().splitlines()
        assert len(rows) == 1 + 13 and rows[0].startswith("frame,body,px")
