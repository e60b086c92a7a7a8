"""GPU parity for the reference's live path: solver::step of one body against the ground
(src/solver.rs:3-27, src/collision.rs:13-35, src/rigid.rs, src/constraint.rs) — configs C1 and C2.
The bar is max position error <= 1e-4 m and orientation error <= 1e-4 rad against the fp32 oracle;
what the kernel is built to deliver, and what these tests also assert, is bit-identical state."""
import numpy as np
import pytest

import oracle_lib as xo
import parity
import constraint_solver_b200 as cs
from constraint_solver_b200 import scenes

pytestmark = pytest.mark.gpu
DT = scenes.FRAME_DT
CUBE = cs.Polytope.new_cube()


def _oracle(rec, frames, substeps=25, verts=CUBE.vertices):
    r = xo.rigid38(rec).astype(np.float32)
    n = xo.step_bodies(r, verts, DT, substeps, frames, threads=xo.hardware_threads())
    return r, n


def test_c1_default_cube_600_frames_every_checkpoint():
    """C1: World::new body a, 25 substeps, checked after N in {1, 10, 60, 600} frames and per frame."""
    a = scenes.cube_body_a().reshape(1, 1)
    cpu = xo.rigid38(a).astype(np.float32)
    with cs.Worlds(1) as w:
        w.add_hull(CUBE)
        w.upload(a)
        total = 0
        for frame in range(1, 601):
            w.integrate(DT, 25, 1)
            total += w.stats()["constraints"]
            n = xo.step_bodies(cpu, CUBE.vertices, DT, 25, 1)
            total -= n
            if frame in (1, 10, 60, 600) or frame % 50 == 0:
                pos, rot, exact = parity.compare(w.download(), cpu)
                assert pos <= parity.POS_TOL and rot <= parity.ROT_TOL, (frame, pos, rot)
                assert exact, f"frame {frame}: GPU state is not bit-identical to the fp32 oracle"
        assert total == 0, "constraint counts differ"


def test_c1_fused_600_frames_one_launch_equals_per_frame_launches():
    a = scenes.cube_body_a().reshape(1, 1)
    cpu, n = _oracle(a, 600)
    with cs.Worlds(1) as w:
        w.add_hull(CUBE); w.upload(a)
        w.integrate(DT, 25, 600)
        assert w.stats()["constraints"] == n and w.stats()["kernel_launches"] == 1
        pos, rot, exact = parity.compare(w.download(), cpu)
        assert exact and pos == 0 and rot == 0


def test_world_b_tetra_metrics_against_cube_hull():
    """world.rs:42 quirk: body b (tetrahedron metrics, off-diagonal inverse inertia) is stepped against p1."""
    wd = scenes.default_world()
    b = wd.b.rec.reshape(1, 1)
    cpu, n = _oracle(b, 120)
    with cs.Worlds(1) as w:
        w.add_hull(CUBE); w.upload(b); w.integrate(DT, 25, 120)
        pos, rot, exact = parity.compare(w.download(), cpu)
        assert exact, (pos, rot)


@pytest.mark.parametrize("frames", [1, 10, 60])
def test_c2_identical_copies_all_equal_c1(frames):
    n = 1 << 14
    cpu, _ = _oracle(scenes.cube_body_a().reshape(1, 1), frames)
    with cs.Worlds(n) as w:
        w.add_hull(CUBE); w.upload_broadcast(scenes.cube_body_a()); w.integrate(DT, 25, frames)
        got = xo.rigid38(w.download())
    assert np.array_equal(got, np.broadcast_to(cpu.astype(np.float64), got.shape))


@pytest.mark.parametrize("frames", [1, 10, 60, 240])
def test_c2_perturbed_sample(frames):
    n = 4096
    init = scenes.c2_perturbed(n)
    cpu, ncons = _oracle(init, frames)
    with cs.Worlds(n) as w:
        w.add_hull(CUBE); w.upload(init); w.integrate(DT, 25, frames)
        pos, rot, exact = parity.compare(w.download(), cpu)
        assert pos <= parity.POS_TOL and rot <= parity.ROT_TOL, (pos, rot)
        assert exact
        assert w.stats()["constraints"] == ncons


@pytest.mark.parametrize("n", [4097, 70_000])
def test_c2_regrouped_bodies_frame_by_frame(n, monkeypatch):
    """Regrouping (xpbd_single.cu): after every step call the thread -> body permutation is rebuilt from the contact
    counts, so from the second call on every body is stepped by a different thread.  Per-frame calls over a settling
    scene (airborne, tumbling and resting bodies mixed), ragged body count; bit-identical to the oracle, and the
    constraint count of every call matches.  n = 70 000 takes the automatic choice (>= 2^16 bodies), 4 097 forces it."""
    if n < (1 << 16):
        monkeypatch.setenv("XPBD_SINGLE_REGROUP", "1")
    frames = 40
    init = scenes.c2_perturbed(n)
    cpu, ncons = _oracle(init, frames)
    total = 0
    with cs.Worlds(n) as w:
        w.add_hull(CUBE); w.upload(init)
        for _ in range(frames):
            w.integrate(DT, 25, 1)
            st = w.stats()
            total += st["constraints"]
            assert st["kernel_launches"] == 4  # step kernel + histogram, scan, scatter
        pos, rot, exact = parity.compare(w.download(), cpu)
    assert exact, (pos, rot)
    assert total == ncons


def test_c2_full_size_properties_and_sampled_parity():
    """C2 at BASELINE size (2^20 worlds): sampled worlds vs the oracle + size-independent properties
    (identical inputs -> identical outputs; unit quaternions; no NaN; nothing below the ground by more than slop)."""
    n = 1 << 20
    init = scenes.c2_perturbed(n)
    init[1::2] = init[0::2]  # every odd world duplicates its even neighbour
    frames = 30
    with cs.Worlds(n) as w:
        w.add_hull(CUBE); w.upload(init); w.integrate(DT, 25, frames)
        out = w.download()
        assert w.stats()["body_substeps"] == n * 25 * frames and w.stats()["nan_worlds"] == 0
    g = xo.rigid38(out)
    assert np.isfinite(g).all()
    assert np.array_equal(g[0::2], g[1::2])
    np.testing.assert_allclose(np.linalg.norm(g[:, 34:38], axis=1), 1.0, atol=1e-5)
    idx = np.arange(0, n, 1021)
    cpu, _ = _oracle(init[idx], frames)
    assert np.array_equal(g[idx], cpu.astype(np.float64))


def test_edge_cases_ragged_ranges_and_other_hulls():
    """Sub-range upload/download, non-multiple-of-block sizes, per-body hull ids, substep counts != 25."""
    n = 1000 + 37
    init = scenes.c2_perturbed(n)
    ico = cs.Polytope.new_icosahedron()
    tet = cs.Polytope.new_tetrahedron()
    ids = (np.arange(n) % 3).astype(np.int32)
    hulls = [CUBE, ico, tet]
    with cs.Worlds(n) as w:
        for h in hulls:
            w.add_hull(h)
        w.set_hulls(ids)
        w.upload(init[:500], 0); w.upload(init[500:], 500)
        w.integrate(DT, 7, 90)
        out = np.concatenate([w.download(0, 123), w.download(123, n - 123)])
    g = xo.rigid38(out)
    for k, h in enumerate(hulls):
        sel = np.where(ids == k)[0]
        cpu, _ = _oracle(init[sel], 90, substeps=7, verts=h.vertices)
        assert np.array_equal(g[sel], cpu.astype(np.float64)), f"hull {k}"


def test_nan_world_does_not_poison_neighbours():
    n = 64
    init = scenes.c2_perturbed(n)
    init["position"][5, 0, 2] = np.nan
    cpu, _ = _oracle(init, 20)
    with cs.Worlds(n) as w:
        w.add_hull(CUBE); w.upload(init); w.integrate(DT, 25, 20)
        assert w.stats()["nan_worlds"] == 1  # failure detection: the caller learns it without downloading the state
        g = xo.rigid38(w.download())
    ok = np.arange(n) != 5
    assert np.array_equal(g[ok], cpu.astype(np.float64)[ok])
    assert np.isnan(g[5, 31:38]).any()


def test_step_host_rigid_single_body_path():
    """xpbd_batch_step_host_rigid on the reference's live path (one body per world): f64 records in and out."""
    n = 300
    init = scenes.c2_perturbed(n)
    cpu, _ = _oracle(init, 12)
    with cs.Worlds(n) as w:
        w.add_hull(CUBE)
        rec = init.copy()
        for _ in range(12):
            rec = w.integrate_host_rigid(rec, DT, 25, 1)
    got, dyn, const = xo.rigid38(rec), np.r_[22:28, 31:38], np.r_[0:22, 28:31]
    assert np.array_equal(got[:, dyn], cpu.astype(np.float64)[:, dyn])
    assert np.array_equal(got[:, const], xo.rigid38(init)[:, const])  # the caller's f64 constants, not their fp32 roundings


def test_dynamic_snapshot_roundtrip_and_rewind():
    """The reference's rewind buffer (src/app.rs:48,204-215): snapshot, step, restore, step again -> same state."""
    n = 256
    init = scenes.c2_perturbed(n)
    with cs.Worlds(n) as w:
        w.add_hull(CUBE); w.upload(init)
        w.integrate(DT, 25, 20)
        snap = w.download_dynamic()
        w.integrate(DT, 25, 20)
        first = xo.rigid38(w.download())
        w.upload_dynamic(snap); w.sync()
        w.integrate(DT, 25, 20)
        again = xo.rigid38(w.download())
    assert np.array_equal(first, again)
    assert snap.shape == (n, 1, 13)
