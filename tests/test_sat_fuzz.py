"""GPU-side A/B fuzz of the SAT stage (VERDICT r01 "Next round" 1a): every schedule of the stage (warp per candidate,
two candidates per warp, thread per candidate), with and without the witness cull, must return the same (feature,
index) for every candidate as the reference's full edge-pair scan (src/collision.rs:151-197 as restated by the
warp-per-candidate kernel without the cull, which in turn is checked against the CPU oracle on a sample).

>= 1e7 synthetic candidates through the step's own kernels (xpbd_sat_feature_batch): unit cubes, boxes scaled 0.1-10
per axis, tetrahedron / icosahedron / QuickHull pairs; offsets 0 / 1e3 / 1e4 m; near-parallel edges (1e-7 ... 1e-3 rad);
resting stacks with penetrations swept ulp by ulp through 0; crossing edges whose gap sweeps through 0 and through the
feature threshold T = min(0, 0.9 face_max + 0.0025)."""
import numpy as np
import pytest

import oracle_world as xw
import constraint_solver_b200 as cs
from constraint_solver_b200 import _abi, collision, scenes

pytestmark = pytest.mark.gpu

CUBE = cs.Polytope.new_cube()
N = 400_000  # candidates per regime call


def box(sx, sy, sz):
    """Cube-shaped hull (8 vertices, 12 edges, 6 quads) scaled per axis."""
    s = np.array([sx, sy, sz], dtype=np.float64)
    return cs.Polytope(CUBE.vertices * s, CUBE.edges, CUBE.faces, CUBE.centroid * s)


# ---- quaternion helpers (v.x v.y v.z s layout) ------------------------------------------------------------------
def qmul(a, b):
    av, as_, bv, bs = a[:, :3], a[:, 3:], b[:, :3], b[:, 3:]
    v = as_ * bv + bs * av + np.cross(av, bv)
    s = as_ * bs - np.sum(av * bv, axis=1, keepdims=True)
    return np.concatenate([v, s], axis=1)


def qrot(q, v):
    qv, s = q[:, :3], q[:, 3:]
    t = np.cross(qv, v) + v * s
    return np.cross(qv, t) * 2.0 + v


def axis_angle(axis, angle):
    axis = axis / np.linalg.norm(axis, axis=1, keepdims=True)
    return np.concatenate([axis * np.sin(0.5 * angle)[:, None], np.cos(0.5 * angle)[:, None]], axis=1)


def small_rotations(rng, n, lo=-7.0, hi=-3.0):
    return axis_angle(rng.normal(size=(n, 3)), 10.0 ** rng.uniform(lo, hi, n))


def cube_group(rng, n):
    """Random elements of the rotation group of the cube (compositions of quarter turns)."""
    q = np.tile(np.array([[0.0, 0.0, 0.0, 1.0]]), (n, 1))
    eye = np.eye(3)
    for _ in range(3):
        ax = eye[rng.integers(0, 3, n)]
        q = qmul(axis_angle(ax, 0.5 * np.pi * rng.integers(0, 4, n)), q)
    return q


def mat_to_quat(R):
    """Rotation matrices (n, 3, 3) -> quaternions (v, s); numerically safe branch per row."""
    n = len(R)
    q = np.zeros((n, 4))
    tr = R[:, 0, 0] + R[:, 1, 1] + R[:, 2, 2]
    # vectorised Shepperd
    c = np.stack([tr, R[:, 0, 0], R[:, 1, 1], R[:, 2, 2]], axis=1).argmax(axis=1)
    for k in range(4):
        m = c == k
        if not m.any():
            continue
        Rm = R[m]
        if k == 0:
            s = np.sqrt(1.0 + tr[m]) * 2
            q[m, 3] = 0.25 * s
            q[m, 0] = (Rm[:, 2, 1] - Rm[:, 1, 2]) / s; q[m, 1] = (Rm[:, 0, 2] - Rm[:, 2, 0]) / s; q[m, 2] = (Rm[:, 1, 0] - Rm[:, 0, 1]) / s
        else:
            i = k - 1; j = (i + 1) % 3; l = (i + 2) % 3
            s = np.sqrt(1.0 + Rm[:, i, i] - Rm[:, j, j] - Rm[:, l, l]) * 2
            q[m, i] = 0.25 * s
            q[m, 3] = (Rm[:, l, j] - Rm[:, j, l]) / s
            q[m, j] = (Rm[:, j, i] + Rm[:, i, j]) / s; q[m, l] = (Rm[:, l, i] + Rm[:, i, l]) / s
    return q / np.linalg.norm(q, axis=1, keepdims=True)


def frames(q, centre, hull):
    """Frame records that put the hull's centroid at `centre` with orientation q (Frame * v = q v + p)."""
    f = np.zeros(len(q), dtype=_abi.FRAME_DTYPE)
    f["rotation"] = q
    f["position"] = centre - qrot(q, np.broadcast_to(hull.centroid, (len(q), 3)))
    return f


def ulp_sweep(rng, n, centre, scale):
    """A mix of uniform values in centre +- scale and values stepping one fp32 ulp at a time through `centre`."""
    out = centre + rng.uniform(-scale, scale, n)
    k = n // 4
    step = np.spacing(np.float32(1.0 + abs(centre)))  # ulp of coordinates of that size
    out[:k] = centre + step * rng.integers(-200, 201, k)
    return out


# ---- regimes ------------------------------------------------------------------------------------------------------
def regime_random(rng, n, ha, hb, scale):
    qa, qb = scenes.random_unit_quaternions(rng, n), scenes.random_unit_quaternions(rng, n)
    ca = rng.uniform(-0.2, 0.2, (n, 3))
    d = rng.normal(size=(n, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
    cb = ca + d * (rng.uniform(0.45, 1.75, n) * scale)[:, None]
    return frames(qa, ca, ha), frames(qb, cb, hb)


def regime_stack(rng, n, ha, hb, height):
    """B resting on A (C3's contact): tiny rotations, x/y jitter, penetration swept through 0 ulp by ulp."""
    qa, qb = small_rotations(rng, n, -7, -2), small_rotations(rng, n, -7, -2)
    ca = np.stack([rng.uniform(-0.01, 0.01, n), rng.uniform(-0.01, 0.01, n), rng.uniform(0.5, 9.5, n)], axis=1)
    dz = ulp_sweep(rng, n, height, 0.01)
    cb = ca + np.stack([rng.uniform(-0.02, 0.02, n), rng.uniform(-0.02, 0.02, n), dz], axis=1)
    return frames(qa, ca, ha), frames(qb, cb, hb)


def regime_parallel(rng, n, ha, hb, height):
    """Arbitrary orientation of A; B = A turned by an element of the cube group and by 1e-7 ... 1e-3 rad: families of
    nearly parallel edges (cross products that are rounding noise), in face contact along a random local axis of A."""
    qa = scenes.random_unit_quaternions(rng, n)
    qb = qmul(qa, qmul(small_rotations(rng, n, -7, -3), cube_group(rng, n)))
    ca = rng.uniform(-0.5, 0.5, (n, 3))
    local = np.eye(3)[rng.integers(0, 3, n)] * rng.choice([-1.0, 1.0], n)[:, None]
    tang = rng.uniform(-0.3, 0.3, (n, 3)) * (1.0 - np.abs(local))
    off = local * ulp_sweep(rng, n, height, 0.02)[:, None] + tang
    cb = ca + qrot(qa, off)
    return frames(qa, ca, ha), frames(qb, cb, hb)


def regime_crossing_edges(rng, n, ha, hb):
    """Unit cubes: an edge of B crosses A's edge (x = 1, z = 1) from the diagonal direction; the gap sweeps through 0 and
    through the edge / face feature threshold; crossing angle from 1e-7 rad (nearly parallel) to pi/2."""
    qa = small_rotations(rng, n, -8, -4)
    ca = np.tile(np.array([[0.5, 0.5, 0.5]]), (n, 1)) + rng.uniform(-1e-3, 1e-3, (n, 3))
    nrm = np.array([1.0, 0.0, 1.0]) / np.sqrt(2.0)
    phi = rng.uniform(0.15, np.pi - 0.15, n)
    k = n // 5
    phi[:k] = 10.0 ** rng.uniform(-7, -1, k)
    t = np.cos(phi)[:, None] * np.array([0.0, 1.0, 0.0]) + np.sin(phi)[:, None] * (np.array([-1.0, 0.0, 1.0]) / np.sqrt(2.0))
    f1, f2 = t, np.broadcast_to(nrm, (n, 3))
    f3 = np.cross(f1, f2)
    e1, e2 = np.array([1.0, 0.0, 0.0]), np.array([0.0, 1.0, 1.0]) / np.sqrt(2.0)
    e3 = np.cross(e1, e2)
    R = f1[:, :, None] * e1[None, None, :] + f2[:, :, None] * e2[None, None, :] + f3[:, :, None] * e3[None, None, :]
    qb = mat_to_quat(R)
    gap = ulp_sweep(rng, n, 0.0, 0.012)
    gap[n // 2:] = ulp_sweep(rng, n - n // 2, -0.0025 / 0.1, 0.02) * 0.1  # around the threshold for shallow face depths
    mid = np.array([1.0, 0.5, 1.0]) + nrm * gap[:, None] + t * rng.uniform(-0.3, 0.3, n)[:, None]
    pb = mid - qrot(qb, np.broadcast_to(np.array([0.5, 0.0, 0.0]), (n, 3)))
    fb = np.zeros(n, dtype=_abi.FRAME_DTYPE)
    fb["rotation"], fb["position"] = qb, pb
    return frames(qa, ca, ha), fb


def shift(fa, fb, offset):
    fa = fa.copy(); fb = fb.copy()
    fa["position"] += offset; fb["position"] += offset
    return fa, fb


def _ab(fa, fb, ha, hb, schedules, oracle_sample=0, seed=0):
    """Baseline = warp per candidate, full scan; every other (schedule, cull) must agree on every candidate."""
    base_f, base_i = collision.sat_feature_batch(fa, fb, ha, hb, "warp", cull=False)
    for sched in schedules:
        for cull in (False, True):
            if sched == "warp" and not cull:
                continue
            f, i = collision.sat_feature_batch(fa, fb, ha, hb, sched, cull=cull)
            bad = np.flatnonzero((f != base_f) | (i != base_i))
            assert bad.size == 0, (sched, cull, bad[:5], f[bad[:5]], base_f[bad[:5]], i[bad[:5]], base_i[bad[:5]])
    if oracle_sample:
        rng = np.random.default_rng(seed)
        idx = rng.choice(len(fa), oracle_sample, replace=False)
        f7 = lambda fr: np.concatenate([fr["position"], fr["rotation"]], axis=1)
        a7, b7 = f7(fa[idx]), f7(fb[idx])
        for k, i in enumerate(idx):
            _, _, feature, _ = xw.sat_manifold(a7[k], b7[k], ha, hb, np.float32)
            assert feature == base_f[i], (i, feature, base_f[i])
    return base_f


CUBE_SCHEDULES = ("warp", "dual", "thread")
OFFSETS = [(0.0, 0.0, 0.0), (1000.0, -1000.0, 1000.0), (-1e4, 1e4, 3e3)]


@pytest.mark.parametrize("offset", OFFSETS, ids=["origin", "1e3", "1e4"])
def test_cube_pairs_every_schedule_and_cull_agree(offset):
    """5 regimes x 400 k unit-cube candidates per offset = 2 M candidates x 3 offsets, each through 6 kernel variants."""
    rng = np.random.default_rng(int(abs(offset[0])) + 17)
    off = np.array(offset)
    seen = set()
    regimes = [regime_random(rng, N, CUBE, CUBE, 1.0), regime_stack(rng, N, CUBE, CUBE, 1.0),
               regime_parallel(rng, N, CUBE, CUBE, 1.0), regime_crossing_edges(rng, N, CUBE, CUBE),
               regime_stack(rng, N, CUBE, CUBE, 0.9999)]
    for fa, fb in regimes:
        fa, fb = shift(fa, fb, off)
        f = _ab(fa, fb, CUBE, CUBE, CUBE_SCHEDULES, oracle_sample=1500 if offset[0] == 0.0 else 300)
        seen.update(np.unique(f).tolist())
    if offset[0] == 0.0:
        assert seen == {0, 1, 2, 3}, seen  # separated, both face features, edge-edge


@pytest.mark.parametrize("sa,sb", [((0.1, 0.1, 0.1), (1.0, 1.0, 1.0)), ((10.0, 10.0, 10.0), (1.0, 1.0, 1.0)),
                                   ((2.0, 0.5, 1.0), (0.3, 3.0, 1.0)), ((10.0, 0.1, 3.0), (10.0, 0.1, 3.0))],
                         ids=["0.1-1", "10-1", "slabs", "plates"])
def test_scaled_boxes_every_schedule_and_cull_agree(sa, sb):
    """Boxes scaled 0.1 ... 10 per axis (still cube-shaped: all three schedules), 3 regimes x 400 k, two offsets."""
    ha, hb = box(*sa), box(*sb)
    rng = np.random.default_rng(101)
    scale = 0.5 * (np.linalg.norm(sa) + np.linalg.norm(sb)) / np.sqrt(3.0)
    height = 0.5 * (sa[2] + sb[2])
    for off in (np.zeros(3), np.array([1000.0, 1000.0, -1000.0])):
        for fa, fb in (regime_random(rng, N, ha, hb, scale), regime_stack(rng, N, ha, hb, height),
                       regime_stack(rng, N, ha, hb, height - 1e-4)):
            fa, fb = shift(fa, fb, off)
            _ab(fa, fb, ha, hb, CUBE_SCHEDULES, oracle_sample=200)


def test_hexahedra_that_are_not_boxes_take_the_per_pair_cull():
    """A frustum has 8 vertices, 12 edges and 6 quads — every all-cube schedule applies — but its slanted edges are not
    parallel: the hull table must not classify it as box-like (HullDesc::box_classes), and the thread-per-candidate
    kernel's class-level cull must fall back to the per-pair tests.  Frustum vs frustum and frustum vs cube."""
    v = CUBE.vertices.copy()
    top = v[:, 2] == 1.0
    v[top, 0:2] = 0.5 + 0.6 * (v[top, 0:2] - 0.5)
    frustum = cs.Polytope(v, CUBE.edges, CUBE.faces, (0.5, 0.5, 0.45))
    rng = np.random.default_rng(23)
    for ha, hb in ((frustum, frustum), (CUBE, frustum), (frustum, CUBE)):
        for fa, fb in (regime_random(rng, N // 2, ha, hb, 1.0), regime_stack(rng, N // 2, ha, hb, 1.0)):
            _ab(fa, fb, ha, hb, CUBE_SCHEDULES, oracle_sample=200)


def test_general_hull_pairs_cull_agrees_with_full_scan():
    """Tetrahedron / icosahedron / QuickHull hulls: the warp-per-candidate kernel with and without the cull."""
    hulls = scenes.c4_hulls()
    rng = np.random.default_rng(7)
    pairs = [(1, 2), (2, 2), (0, 5), (3, 7), (6, 4), (7, 7), (1, 1)]
    n = 150_000
    for ia, ib in pairs:
        ha, hb = hulls[ia], hulls[ib]
        fa, fb = regime_random(rng, n, ha, hb, 0.7)
        fa2, fb2 = shift(fa, fb, np.array([0.0, 2000.0, 0.0]))
        _ab(fa, fb, ha, hb, ("warp",), oracle_sample=150)
        _ab(fa2, fb2, ha, hb, ("warp",), oracle_sample=50)


def _long_run(n, frames, env, monkeypatch):
    for k in ("XPBD_SAT_SCHED", "XPBD_SAT_CULL", "XPBD_SAT_ORDER", "XPBD_WORLD_PATH"):
        monkeypatch.delenv(k, raising=False)
    monkeypatch.setenv("XPBD_WORLD_PATH", "staged")
    for k, v in env.items():
        monkeypatch.setenv(k, v)  # read when the batch builds its staged state / at every step call
    with cs.Worlds(n, 10, device=0, body_body_contacts=True) as w:
        w.add_hull(CUBE)
        w.upload(scenes.c3_stack(n))
        w.integrate(scenes.FRAME_DT, 20, frames)
        st = w.stats()
        assert st["overflow_worlds"] == 0 and st["nan_worlds"] == 0
        return w.download_dynamic().copy(), st["constraints"]


@pytest.mark.parametrize("variant", [{"XPBD_SAT_CULL": "0"}, {"XPBD_SAT_ORDER": "list"}, {"XPBD_SAT_SCHED": "dual"},
                                     {"XPBD_SAT_SCHED": "warp"}, {"XPBD_SAT_SCHED": "warp", "XPBD_SAT_CULL": "0"}],
                         ids=["thread-nocull", "thread-worklist", "dual", "warp", "warp-nocull"])
def test_whole_scene_long_run_every_sat_variant_same_bits(variant, monkeypatch):
    """The fuzz above checks the SAT stage candidate by candidate; this one lets a difference anywhere in it COMPOUND: 8 192
    ten-cube stacks (the C3 scene: 73 728 resting contacts whose depths sit around the cull's rounding margin) stepped 300
    frames = 6 000 substeps = 4.4e8 candidate evaluations per run, with the default SAT stage (thread per candidate, class-level
    + per-pair witness cull, (pair slot, world) order) and with each variant — the reference's full edge-pair scan, the work
    list, the other schedules.  Chaotic dynamics amplify a single differing feature choice into different states within a few
    frames; the final states and constraint counts must be bit-identical."""
    n, frames = 8192, 300
    base, c0 = _long_run(n, frames, {}, monkeypatch)
    got, c1 = _long_run(n, frames, variant, monkeypatch)
    assert c0 == c1
    assert np.array_equal(base.view(np.uint32), got.view(np.uint32))


def _long_run_c4(n, frames, env, monkeypatch):
    for k in ("XPBD_SAT_SCHED", "XPBD_SAT_CULL", "XPBD_SAT_ORDER", "XPBD_WORLD_PATH", "XPBD_SOLVE_LANES", "XPBD_SUB_BATCHES"):
        monkeypatch.delenv(k, raising=False)
    monkeypatch.setenv("XPBD_WORLD_PATH", "staged")
    for k, v in env.items():
        monkeypatch.setenv(k, v)
    init, hulls, ids, joints = scenes.c4_pile(n)
    with cs.Worlds(n, init.shape[1], device=0, body_body_contacts=True, joints_per_world=joints.shape[1]) as w:
        for h in hulls:
            w.add_hull(h)
        w.set_hulls(ids)
        w.set_joints(joints)
        w.upload(init)
        w.integrate(scenes.FRAME_DT, 20, frames)
        st = w.stats()
        assert st["overflow_worlds"] == 0 and st["nan_worlds"] == 0
        return w.download_dynamic().copy(), st["constraints"]


@pytest.mark.parametrize("variant", [{"XPBD_SAT_CULL": "0"}, {"XPBD_SOLVE_LANES": "2"}, {"XPBD_SOLVE_LANES": "32", "XPBD_SUB_BATCHES": "1"}],
                         ids=["nocull", "solve-lane-pair", "solve-32-lanes-one-range"])
def test_mixed_hull_scene_long_run_every_variant_same_bits(variant, monkeypatch):
    """The same for the C4 scene (1 024 piles of 64 mixed QuickHull hulls + joints, 50 frames: the piles land from frame 19 on
    and the candidate count peaks around frame 34): the warp-per-candidate SAT kernel with and without the witness cull, the
    solve stage as one lane pair per world and as 16 / 32 lanes per world by dependency level — bit-identical final states."""
    n, frames = 1024, 50
    base, c0 = _long_run_c4(n, frames, {}, monkeypatch)
    got, c1 = _long_run_c4(n, frames, variant, monkeypatch)
    assert c0 == c1
    assert np.array_equal(base.view(np.uint32), got.view(np.uint32))
