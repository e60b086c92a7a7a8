"""GPU parity of the multi-body world step against the fp32 oracle
(oracle/xo_world.hpp — the builder's own specification: the reference has no algorithm here).
Bar: <= 1e-4 m / 1e-4 rad; asserted: bit-identical state and identical constraint counts."""
import numpy as np
import pytest

import oracle_lib as xo
import oracle_world as xw
import parity
import constraint_solver_b200 as cs
from constraint_solver_b200 import _abi, scenes

pytestmark = pytest.mark.gpu
DT = scenes.FRAME_DT


@pytest.fixture(autouse=True, params=["staged", "fused"])
def world_path(request, monkeypatch):
    """Every test runs on both multi-body paths: the staged pipeline (xpbd_staged.cu) and the fused
    single-launch kernel (xpbd_world.cu).  The library reads XPBD_WORLD_PATH at each step call."""
    monkeypatch.setenv("XPBD_WORLD_PATH", request.param)
    return request.param

CUBE = cs.Polytope.new_cube()


def _run_gpu(init, hulls, frames, substeps, hull_ids=None, joints=None, body_body=True, per_frame=False, sub_batches=0, **kw):
    n, B = init.shape
    J = 0 if joints is None else joints.shape[1]
    with cs.Worlds(n, B, joints_per_world=J, body_body_contacts=body_body, **kw) as w:
        if sub_batches:
            w.set_sub_batches(sub_batches)
        for h in hulls:
            w.add_hull(h)
        if hull_ids is not None:
            w.set_hulls(hull_ids)
        if joints is not None:
            w.set_joints(joints)
        w.upload(init)
        cons = 0
        if per_frame:
            for _ in range(frames):
                w.integrate(DT, substeps, 1)
                cons += w.stats()["constraints"]
        else:
            w.integrate(DT, substeps, frames)
            cons = w.stats()["constraints"]
        st = w.stats()
        assert st["overflow_worlds"] == 0
        if sub_batches and st["kernel_launches"] > 1:  # the staged pipeline ran: it must have used that many world ranges
            assert st["sub_batches"] == min(sub_batches, n)
        return w.download(), cons


def _run_cpu(init, hulls, frames, substeps, hull_ids=None, joints=None, body_body=True, **kw):
    n, B = init.shape
    r = xo.rigid38(init).astype(np.float32)
    c = xw.step_worlds(r, B, list(hulls), DT, substeps, frames, threads=xo.hardware_threads(), hull_ids=hull_ids,
                       joints=joints, body_body=body_body, **kw)
    return r, c


def _check(init, hulls, frames, substeps, sub_batches=0, **kw):
    gpu, gcons = _run_gpu(init, hulls, frames, substeps, sub_batches=sub_batches, **kw)
    cpu, c = _run_cpu(init, hulls, frames, substeps, **kw)
    pos, rot, exact = parity.compare(gpu, cpu)
    assert pos <= parity.POS_TOL and rot <= parity.ROT_TOL, (pos, rot)
    assert exact, f"not bit-identical (max pos err {pos}, rot err {rot})"
    assert gcons == c["constraints"], (gcons, c)
    return gpu, c


def test_single_body_world_path_equals_reference_path():
    """B = 1 through the multi-body kernel reduces to solver::step bit for bit (ground constraints are the reference's)."""
    init = scenes.c2_perturbed(512)
    gpu, _ = _run_gpu(init, [CUBE], 60, 25, body_body=True)
    r = xo.rigid38(init).astype(np.float32)
    xo.step_bodies(r, CUBE.vertices, DT, 25, 60, threads=4)
    assert np.array_equal(xo.rigid38(gpu), r.astype(np.float64))


@pytest.mark.parametrize("frames", [1, 10, 60])
def test_c3_stack_parity(frames):
    init = scenes.c3_stack(64)
    gpu, c = _check(init, [CUBE], frames, 20)
    if frames >= 10:
        assert c["manifolds"] > 0


def test_c3_stack_far_from_the_origin():
    """The same stacks 800 m away from the origin: fp32 positions carry ~6e-5 of rounding noise there, which the SAT
    stage's witness cull has to absorb in its margin (it scales with the coordinates) — still bit-identical."""
    init = scenes.c3_stack(48)
    init["position"][:, :, 0] += 731.0
    init["position"][:, :, 1] -= 415.0
    _, c = _check(init, [CUBE], 40, 20)
    assert c["manifolds"] > 0


def test_c3_stack_long_run_stays_a_stack():
    """600 frames: parity on every world + physical sanity."""
    init = scenes.c3_stack(32)
    gpu, _ = _check(init, [CUBE], 600, 20)
    g = xo.rigid38(gpu).reshape(32, 10, 38)
    assert np.isfinite(g).all()
    dz = g[:, :, 33] - np.arange(10)[None, :]
    assert np.abs(dz).max() < 0.02, "cubes left their level in the stack"
    assert np.abs(g[:, :, 31:33]).max() < 0.6, "stack slid apart"
    assert np.abs(g[:, :, 22:28]).max() < 1.0, "stack did not come to rest"


def test_per_frame_launches_equal_fused_launch():
    init = scenes.c3_stack(16)
    a, ca = _run_gpu(init, [CUBE], 12, 20)
    b, cb = _run_gpu(init, [CUBE], 12, 20, per_frame=True)
    assert np.array_equal(xo.rigid38(a), xo.rigid38(b)) and ca == cb


def test_tumbling_pile_of_mixed_hulls():
    """Random poses: exercises separated pairs, all three contact features and face/edge ties."""
    hulls = [CUBE, cs.Polytope.new_icosahedron().scaled(0.6), cs.Polytope.new_tetrahedron()]
    n, B = 97, 6  # odd: with two worlds per warp in stage_prepare the last half-warp idles (compact vertex enumeration)
    rng = np.random.default_rng(5)
    init = np.empty((n, B), dtype=_abi.RIGID_DTYPE)
    ids = rng.integers(0, 3, (n, B)).astype(np.int32)
    for k, h in enumerate(hulls):
        r = cs.Rigid(h.rigid_metrics(0.5))
        r.external_force = (0.0, 0.0, -2.0 / r.inverse_mass * 0.1)
        init[ids == k] = r.rec
    init["position"][:, :, 0:2] = rng.uniform(-0.6, 0.6, (n, B, 2))
    init["position"][:, :, 2] = 0.3 + 1.1 * np.arange(B)[None, :] + rng.uniform(0, 0.2, (n, B))
    init["rotation"] = scenes.random_unit_quaternions(rng, n * B).reshape(n, B, 4)
    init["angular_velocity"] = rng.uniform(-2, 2, (n, B, 3))
    _, c = _check(init, hulls, 90, 20, hull_ids=ids)
    assert c["manifolds"] > 1000


def test_tumbling_pile_of_cubes_and_boxes():
    """All hulls cube-shaped (a unit cube and a scaled box): on the staged path this is the two-pairs-per-warp SAT
    kernel, with separated candidates, edge-edge contacts, an odd world count and odd candidate counts (one half of
    the last warp idles)."""
    hulls = [CUBE, CUBE.scaled(0.7)]
    n, B = 51, 5
    rng = np.random.default_rng(11)
    init = np.empty((n, B), dtype=_abi.RIGID_DTYPE)
    ids = rng.integers(0, 2, (n, B)).astype(np.int32)
    for k, h in enumerate(hulls):
        r = cs.Rigid(h.rigid_metrics(0.5))
        r.external_force = (0.0, 0.0, -10.0 / r.inverse_mass)
        init[ids == k] = r.rec
    init["position"][:, :, 0:2] = rng.uniform(-0.4, 0.4, (n, B, 2))
    init["position"][:, :, 2] = 0.5 + 1.5 * np.arange(B)[None, :] + rng.uniform(0, 0.2, (n, B))  # no initial overlap
    init["rotation"] = scenes.random_unit_quaternions(rng, n * B).reshape(n, B, 4)
    init["angular_velocity"] = rng.uniform(-2, 2, (n, B, 3))
    _, c = _check(init, hulls, 120, 20, hull_ids=ids)
    assert c["manifolds"] > 5000 and c["pairs_tested"] > c["manifolds"]  # contacts and separated candidates


def test_joints_chain_without_body_contacts():
    """Distance joints (Constraint.distance, src/constraint.rs:9): a hanging chain of cubes, one end pinned to the world."""
    n, B = 48, 5
    rng = np.random.default_rng(8)
    init = np.empty((n, B), dtype=_abi.RIGID_DTYPE)
    base = scenes.cube_rigid(0.1); base.external_force = (0, 0, -2.0)
    init[:] = base.rec
    init["position"][:, :, 0] = 1.5 * np.arange(B)[None, :]
    init["position"][:, :, 2] = 6.0
    init["velocity"] = rng.uniform(-0.5, 0.5, (n, B, 3))
    joints = np.zeros((n, B), dtype=_abi.JOINT_DTYPE)
    joints["body_a"][:, 0] = 0; joints["body_b"][:, 0] = -1
    joints["anchor_a"][:, 0] = (0.5, 0.5, 1.0); joints["anchor_b"][:, 0] = (0.5, 0.5, 7.5); joints["distance"][:, 0] = 0.5
    for k in range(1, B):
        joints["body_a"][:, k] = k - 1; joints["body_b"][:, k] = k
        joints["anchor_a"][:, k] = (1.0, 0.5, 0.5); joints["anchor_b"][:, k] = (0.0, 0.5, 0.5); joints["distance"][:, k] = 0.5
    joints["body_a"][::7, B - 1] = -1  # inactive slots are skipped
    gpu, c = _check(init, [CUBE], 120, 20, joints=joints, body_body=False)
    g = xo.rigid38(gpu).reshape(n, B, 38)
    assert np.isfinite(g).all() and c["constraints"] > 0


def test_stack_with_attachment_joints_and_contacts():
    n, B = 32, 4
    init = scenes.c3_stack(n, n_cubes=B)
    joints = np.zeros((n, 2), dtype=_abi.JOINT_DTYPE)
    joints["body_a"][:, 0] = 0; joints["body_b"][:, 0] = 1
    joints["anchor_a"][:, 0] = (0.5, 0.5, 1.0); joints["anchor_b"][:, 0] = (0.5, 0.5, 0.0); joints["distance"][:, 0] = 0.0
    joints["body_a"][:, 1] = 2; joints["body_b"][:, 1] = 3
    joints["anchor_a"][:, 1] = (0.0, 0.0, 1.0); joints["anchor_b"][:, 1] = (0.0, 0.0, 0.0); joints["distance"][:, 1] = 0.1
    _check(init, [CUBE], 60, 20, joints=joints)


def test_ragged_world_count_and_friction_compliance_params():
    init = scenes.c3_stack(13, n_cubes=3)  # not a multiple of the worlds-per-block
    _check(init, [CUBE], 30, 10, friction=0.5, compliance=1e-5)


@pytest.mark.parametrize("k", [2, 3, 8])
def test_sub_batches_do_not_change_results(world_path, k):
    """xpbd_batch_set_sub_batches: the staged pipeline steps k contiguous world ranges concurrently on k streams (pointer
    offsets into every per-world / per-body array, hull ids and joints included).  Ragged world counts, more ranges
    than some ranges have worlds to fill a block, joints + mixed hulls; bit-identical to the oracle."""
    if world_path == "fused":
        pytest.skip("sub-batches are a schedule of the staged pipeline")
    init, hulls, ids, joints = scenes.c4_pile(13)
    _, c = _check(init, hulls, 30, 20, hull_ids=ids, joints=joints, sub_batches=k)
    assert c["manifolds"] > 0
    init = scenes.c3_stack(37)
    _check(init, [CUBE], 30, 20, sub_batches=k)
    _check(scenes.c3_stack(5, n_cubes=3), [CUBE], 10, 10, sub_batches=8)  # more ranges than worlds


@pytest.mark.parametrize("lanes", [2, 4, 8, 16, 32])
def test_solve_lane_counts_do_not_change_results(world_path, lanes, monkeypatch):
    """XPBD_SOLVE_LANES: the solve stage with `lanes` lanes per world (2 = one lane pair walks the world's constraints in
    order; 4 ... 32 = stage_solve_wide_kernel: ground contacts by body, manifolds and joints by dependency level, spread
    over lanes / 2 lane pairs).  Constraints that share a body keep their order, so every schedule is bit-identical to the
    oracle: a dense mixed-hull pile with joints (many independent manifolds), joint chains through shared bodies, a stack
    (one dependent chain), ragged world counts."""
    if world_path == "fused":
        pytest.skip("a schedule of the staged pipeline")
    monkeypatch.setenv("XPBD_SOLVE_LANES", str(lanes))
    init, hulls, ids, joints = scenes.c4_pile(11, spacing=1.5)  # dense lattice: body-body contacts from the first frames
    _, c = _check(init, hulls, 20, 20, hull_ids=ids, joints=joints)
    assert c["manifolds"] > 0
    # joints that share bodies (a chain 0-1-2-...-7 in index order, then a second pass in reverse order): several levels
    n, B = 7, 8
    init = scenes.c3_stack(n, n_cubes=B)
    jt = np.zeros((n, 2 * (B - 1)), dtype=_abi.JOINT_DTYPE)
    jt["body_a"] = -1
    order = [(i, i + 1) for i in range(B - 1)] + [(i + 1, i) for i in reversed(range(B - 1))]
    for j, (a, b) in enumerate(order):
        jt["body_a"][:, j] = a; jt["body_b"][:, j] = b
        jt["anchor_a"][:, j] = (0.5, 0.5, 0.5); jt["anchor_b"][:, j] = (0.5, 0.5, 0.5); jt["distance"][:, j] = 1.05
    _check(init, [CUBE], 15, 20, joints=jt)
    _check(scenes.c3_stack(37), [CUBE], 20, 20)
    # random joint graphs, different in every world: repeated bodies, repeated pairs, world anchors (body_b = -1), inactive
    # slots — arbitrary dependency DAGs for the level schedule; bodies on a loose lattice (24 per world: the wide default)
    n, B, J = 9, 24, 40
    rng = np.random.default_rng(1234 + lanes)
    init = np.empty((n, B), dtype=_abi.RIGID_DTYPE)
    base = scenes.cube_rigid(0.3); base.external_force = (0, 0, -3.0)
    init[:] = base.rec
    cell = np.stack(np.meshgrid(np.arange(4), np.arange(3), np.arange(2), indexing="ij"), axis=-1).reshape(B, 3)
    init["position"] = cell[None, :, :] * 1.6 + np.array([0.0, 0.0, 0.4]) + rng.uniform(-0.05, 0.05, (n, B, 3))
    init["velocity"] = rng.uniform(-0.3, 0.3, (n, B, 3))
    jt = np.zeros((n, J), dtype=_abi.JOINT_DTYPE)
    a = rng.integers(0, B, (n, J)); b = rng.integers(0, B, (n, J))
    b = np.where(b == a, (b + 1) % B, b)
    b = np.where(rng.random((n, J)) < 0.15, -1, b)            # anchored to the world
    jt["body_a"] = np.where(rng.random((n, J)) < 0.1, -1, a)  # inactive
    jt["body_b"] = b
    jt["anchor_a"] = rng.uniform(0.0, 1.0, (n, J, 3))
    jt["anchor_b"] = np.where((b == -1)[:, :, None], init["position"][np.arange(n)[:, None], a] + rng.uniform(-0.5, 1.5, (n, J, 3)),
                              rng.uniform(0.0, 1.0, (n, J, 3)))
    # rest lengths near the initial anchor distances (centres, roughly): the bodies are tugged about, not pulled into one clump
    pa = init["position"][np.arange(n)[:, None], a] + 0.5
    pb = np.where((b == -1)[:, :, None], jt["anchor_b"], init["position"][np.arange(n)[:, None], np.maximum(b, 0)] + 0.5)
    jt["distance"] = np.linalg.norm(pb - pa, axis=2) * rng.uniform(0.9, 1.1, (n, J))
    gpu, c = _check(init, [CUBE], 12, 20, joints=jt)
    assert np.isfinite(xo.rigid38(gpu)).all() and c["constraints"] > 0


def test_sub_batches_back_to_back_step_calls(world_path):
    """Between back-to-back step calls the world ranges do not wait for each other (no per-frame barrier); an
    asynchronous upload in between makes them wait for the batch stream again.  Same bits as the oracle either way."""
    if world_path == "fused":
        pytest.skip("sub-batches are a schedule of the staged pipeline")
    n = 257
    init = scenes.c3_stack(n)
    cpu, _ = _run_cpu(init, [CUBE], 12, 20)
    with cs.Worlds(n, 10, body_body_contacts=True) as w:
        w.add_hull(CUBE)
        w.set_sub_batches(4)
        w.upload(init)
        for _ in range(6):
            w.integrate(DT, 20, 1, sync=False)
        snap = w.download_dynamic()          # joins, reads
        w.upload_dynamic(snap)               # asynchronous on the batch stream: the ranges must wait for it
        for _ in range(6):
            w.integrate(DT, 20, 1, sync=False)
        got = w.download()
        assert w.stats()["sub_batches"] == 4
    assert np.array_equal(xo.rigid38(got), cpu.astype(np.float64))


@pytest.mark.parametrize("k", [1, 4])
def test_step_host_equals_upload_step_download(world_path, k):
    """xpbd_batch_step_host_f32 (state in host memory: upload + step + download in one call, the copies pipelined per
    world range) gives the same bits as the three separate calls, frame after frame."""
    n = 131
    init = scenes.c3_stack(n)
    with cs.Worlds(n, 10, body_body_contacts=True) as a, cs.Worlds(n, 10, body_body_contacts=True) as b:
        for w in (a, b):
            w.add_hull(CUBE)
            w.set_sub_batches(k)
            w.upload(init)
        dyn = a.download_dynamic()
        ref = dyn.copy()
        for _ in range(8):
            b.upload_dynamic(ref)
            b.integrate(DT, 20, 1)
            ref = b.download_dynamic()
            dyn = a.integrate_host(dyn, DT, 20, 1)
            assert np.array_equal(dyn, ref)
        # NULL input keeps the device state; NULL output only steps
        a.integrate_host_ptr(DT, 20, 1, 0, 0)
        b.integrate(DT, 20, 1)
        assert np.array_equal(a.download_dynamic(), b.download_dynamic())
    assert np.abs(ref).max() > 0


@pytest.mark.parametrize("k", [1, 4])
def test_step_host_rigid_keeps_f64_constants_and_equals_separate_calls(world_path, k):
    """xpbd_batch_step_host_rigid: the caller's own f64 records (304 B per body) in, the same records out — the 13 dynamic
    values as upload + step + download gives them, every constant field bit for bit as the caller wrote it (not fp32-rounded),
    frame after frame, with `out` aliasing `in`."""
    n = 131
    init = scenes.c3_stack(n)
    init["inverse_inertia"] *= 1.0 + 1e-9   # no longer representable in fp32: a round trip through the device must not round it
    init["center_of_mass"] += 1e-12
    with cs.Worlds(n, 10, body_body_contacts=True) as a, cs.Worlds(n, 10, body_body_contacts=True) as b:
        for w in (a, b):
            w.add_hull(CUBE)
            w.set_sub_batches(k)
        rec = init.copy()
        b.upload(init)
        for _ in range(6):
            rec = a.integrate_host_rigid(rec, DT, 20, 1)
            b.integrate(DT, 20, 1)
            want = xo.rigid38(b.download()).reshape(-1, 38)
            got = xo.rigid38(rec).reshape(-1, 38)
            dyn = np.r_[22:28, 31:38]
            const = np.r_[0:22, 28:31]
            assert np.array_equal(got[:, dyn], want[:, dyn])
            assert np.array_equal(got[:, const], xo.rigid38(init).reshape(-1, 38)[:, const])
        # in place, and NULL input = keep the device state
        a.integrate_host_rigid_ptr(DT, 20, 1, rec.ctypes.data, rec.ctypes.data)
        a.integrate_host_rigid_ptr(DT, 20, 1, 0, rec.ctypes.data)
        b.integrate(DT, 20, 2)
        assert np.array_equal(xo.rigid38(rec).reshape(-1, 38)[:, dyn], xo.rigid38(b.download()).reshape(-1, 38)[:, dyn])


def test_sub_batches_argument_is_validated():
    with cs.Worlds(8, 2, body_body_contacts=True) as w:
        for bad in (-1, 9):
            with pytest.raises(_abi.XpbdError):
                w.set_sub_batches(bad)
        w.set_sub_batches(0)


def test_sub_batches_stats_sum_over_ranges(world_path):
    """constraints / candidate_pairs of xpbd_step_stats are totals over all world ranges."""
    if world_path == "fused":
        pytest.skip("sub-batches are a schedule of the staged pipeline")
    init = scenes.c3_stack(64)
    res = []
    for k in (1, 4):
        with cs.Worlds(64, 10, body_body_contacts=True) as w:
            w.add_hull(CUBE)
            w.set_sub_batches(k)
            w.upload(init)
            w.integrate(DT, 20, 15)
            st = w.stats()
            res.append((st["constraints"], st["candidate_pairs"], st["overflow_worlds"]))
            assert st["sub_batches"] == k and st["kernel_launches"] == k * 4 * 20 * 15 + 1  # + the NaN-world count
    assert res[0] == res[1] and res[0][1] > 0


def test_nan_worlds_are_counted_and_isolated(world_path):
    """xpbd_step_stats.nan_worlds (failure detection): worlds 3 and 40 start with a NaN / Inf; they are
    counted, and every other world still matches the oracle bit for bit."""
    n = 64
    init = scenes.c3_stack(n)
    init["position"][3, 4, 1] = np.nan
    init["velocity"][40, 0, 2] = np.inf
    with cs.Worlds(n, 10, body_body_contacts=True) as w:
        w.add_hull(CUBE); w.upload(init)
        w.integrate(DT, 20, 5)
        st = w.stats()
        got = xo.rigid38(w.download()).reshape(n, 10, 38)
    assert st["nan_worlds"] == 2, st
    init_ok = init.copy()
    init_ok[3], init_ok[40] = init_ok[0], init_ok[0]
    cpu, _ = _run_cpu(init_ok, [CUBE], 5, 20)
    ok = np.ones(n, dtype=bool); ok[[3, 40]] = False
    assert np.array_equal(got[ok], cpu.astype(np.float64).reshape(n, 10, 38)[ok])


def test_default_selector_small_batch_of_large_worlds(monkeypatch):
    """with XPBD_WORLD_PATH unset, a few worlds of many bodies (125 per world do not fit the fused
    kernel's shared memory) must fall back to the staged pipeline instead of failing with XPBD_ERR_CAPACITY."""
    monkeypatch.delenv("XPBD_WORLD_PATH", raising=False)
    init, hulls, ids, joints = scenes.c4_pile(3, n_side=5, spacing=1.5)
    gpu, gcons = _run_gpu(init, hulls, 6, 20, hull_ids=ids, joints=joints)
    cpu, c = _run_cpu(init, hulls, 6, 20, hull_ids=ids, joints=joints)
    assert parity.compare(gpu, cpu)[2] and gcons == c["constraints"]


def _sample_blocks(n, n_blocks, block):
    """Sampled worlds: n_blocks contiguous blocks of `block` worlds spread over [0, n) — first and last worlds, and
    (with 4 automatic world ranges) blocks on both sides of every range boundary are among them."""
    starts = np.linspace(0, n - block, n_blocks).astype(np.int64)
    starts[1:-1] = (starts[1:-1] // 2) * 2  # keep the even/odd duplicate pairs together
    return starts


def _full_size_checkpoints(init, hulls, substeps, checkpoints, starts, block, hull_ids=None, joints=None):
    """Steps the whole batch on the GPU and the sampled worlds on the oracle, comparing at every checkpoint (frames)."""
    n, B = init.shape
    sel = np.concatenate([np.arange(s, s + block) for s in starts])
    cpu = xo.rigid38(init[sel]).astype(np.float32)
    kw = dict(hull_ids=None if hull_ids is None else hull_ids[sel], joints=None if joints is None else joints[sel])
    J = 0 if joints is None else joints.shape[1]
    with cs.Worlds(n, B, joints_per_world=J, body_body_contacts=True) as w:
        for h in hulls:
            w.add_hull(h)
        if hull_ids is not None:
            w.set_hulls(hull_ids)
        if joints is not None:
            w.set_joints(joints)
        w.upload(init)
        done = 0
        for frames in checkpoints:
            w.integrate(DT, substeps, frames - done)
            st = w.stats()
            assert st["overflow_worlds"] == 0 and st["nan_worlds"] == 0, (frames, st)
            xw.step_worlds(cpu, B, list(hulls), DT, substeps, frames - done, threads=xo.hardware_threads(), **kw)
            done = frames
            got = np.concatenate([xo.rigid38(w.download(int(s), block)) for s in starts])
            pos, rot, exact = parity.compare_arrays(got, cpu)
            assert pos <= parity.POS_TOL and rot <= parity.ROT_TOL, (frames, pos, rot)
            assert exact, f"frame {frames}: sampled worlds not bit-identical (max pos err {pos}, rot err {rot})"
        return xo.rigid38(w.download()).reshape(n, B, 38)


def test_c3_full_size_properties_and_sampled_parity(world_path):
    """C3 at BASELINE size (100k worlds x 10 cubes, default schedule): 1 024 sampled worlds (8 blocks of 128) compared
    with the oracle after 1, 10, 60 and 600 frames; duplicated worlds give identical results,
    no NaN, no overflow, stacks stay stacks.  Oracle time: 1 024 worlds x 600 frames = 1.2e8 body-substeps, ~2 min."""
    if world_path == "fused":
        pytest.skip("full-size C3 is the staged pipeline's case; the fused kernel is covered at reduced size")
    n = 100_000
    init = scenes.c3_stack(n)
    init[1::2] = init[0::2]
    g = _full_size_checkpoints(init, [CUBE], 20, (1, 10, 60, 600), _sample_blocks(n, 8, 128), 128)
    assert np.isfinite(g).all()
    assert np.array_equal(g[0::2], g[1::2])
    assert np.abs(g[:, :, 33] - np.arange(10)[None, :]).max() < 0.02


def test_c4_pile_mixed_quickhull_hulls_with_joints():
    """C4 (BASELINE.json configs[3]) at reduced world count: 64 bodies per world drawn from the reference's built-ins
    and QuickHull hulls, 32 distance joints per world; every world compared with the oracle."""
    init, hulls, ids, joints = scenes.c4_pile(24)
    gpu, c = _check(init, hulls, 30, 20, hull_ids=ids, joints=joints)
    g = xo.rigid38(gpu)
    assert np.isfinite(g).all()
    assert c["manifolds"] > 0 and c["constraints"] > 24 * 32 * 30 * 20  # joints alone give that many
    assert g.reshape(24, 64, 38)[:, :, 33].min() > -0.5  # nothing fell through the ground


def test_c4_shard_equals_slice_of_the_whole():
    """Sharding by world index: stepping worlds [8, 16) alone gives the same bits as stepping
    them inside the whole batch."""
    init, hulls, ids, joints = scenes.c4_pile(16)
    whole, _ = _run_gpu(init, hulls, 6, 20, hull_ids=ids, joints=joints)
    init2, _, ids2, joints2 = scenes.c4_pile(8, first_world=8)
    part, _ = _run_gpu(init2, hulls, 6, 20, hull_ids=ids2, joints=joints2)
    assert np.array_equal(xo.rigid38(whole[8:]), xo.rigid38(part))


def test_large_world_without_pair_table(world_path):
    """125 bodies per world: more body pairs than the prepare stage's shared-memory pair table holds (B > 91), so the
    broadphase takes its row-by-row path; dense lattice, so the pair capacity (4 B) is also put under pressure."""
    if world_path == "fused":
        pytest.skip("125 bodies per world exceed the fused kernel's shared-memory world size")
    init, hulls, ids, joints = scenes.c4_pile(3, n_side=5, spacing=1.5)
    _, c = _check(init, hulls, 40, 20, hull_ids=ids, joints=joints)
    assert c["pairs_tested"] > 100_000 and c["manifolds"] > 3000


def test_c4_full_size_properties_and_sampled_parity(world_path):
    """C4 at BASELINE size (16 384 worlds x 64 bodies, 32 joints): 256 sampled worlds (8 blocks of 32) compared with the
    oracle after 1, 10, 60 and 200 frames; duplicated worlds give identical results, no NaN, no capacity overflow,
    nothing tunnels through the ground.  Oracle time: 256 worlds x 200 frames = 6.6e7 body-substeps, ~10 s."""
    if world_path == "fused":
        pytest.skip("full-size C4 is the staged pipeline's case; the fused kernel is covered at reduced size")
    n = 16_384
    init, hulls, ids, joints = scenes.c4_pile(n)
    init[1::2], ids[1::2], joints[1::2] = init[0::2], ids[0::2], joints[0::2]
    g = _full_size_checkpoints(init, hulls, 20, (1, 10, 60, 200), _sample_blocks(n, 8, 32), 32, hull_ids=ids, joints=joints)
    assert np.isfinite(g).all()
    assert np.array_equal(g[0::2], g[1::2])
    assert g[:, :, 33].min() > -0.5
