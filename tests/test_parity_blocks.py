"""GPU parity of every building block behind the per-function C-ABI entry points,
each against the fp32 oracle on the same seeded inputs, bit for bit."""
import ctypes as C

import numpy as np
import pytest

import oracle_lib as xo
import oracle_world as xw
import constraint_solver_b200 as cs
from constraint_solver_b200 import _abi, scenes

pytestmark = pytest.mark.gpu
L = _abi.load
HULLS = {"tetra": (0, cs.Polytope.new_tetrahedron()), "cube": (1, cs.Polytope.new_cube()), "icosa": (2, cs.Polytope.new_icosahedron())}


def _bodies(n, seed=1):
    """Random tumbling bodies near the ground, with off-diagonal inertia and non-zero internal force/torque."""
    rng = np.random.default_rng(seed)
    rec = scenes.c2_perturbed(n).reshape(n)
    rec["position"][:, 2] = rng.uniform(-0.3, 1.0, n)
    b = scenes.default_world().b.rec
    rec["inverse_inertia"][n // 2:] = b["inverse_inertia"]
    rec["center_of_mass"][n // 2:] = b["center_of_mass"]
    rec["internal_force"] = rng.uniform(-1, 1, (n, 3))
    rec["internal_torque"] = rng.uniform(-0.1, 0.1, (n, 3))
    rec["external_torque"] = rng.uniform(-0.1, 0.1, (n, 3))
    return rec


def _frames(n, seed, spread=1.2):
    rng = np.random.default_rng(seed)
    f = np.zeros(n, dtype=_abi.FRAME_DTYPE)
    f["position"] = rng.uniform(-spread, spread, (n, 3))
    f["rotation"] = scenes.random_unit_quaternions(rng, n)
    return f


def _f7(fr):
    return np.concatenate([fr["position"], fr["rotation"]], axis=-1)


def test_integrate_derive_frame_apply_impulse():
    n = 513
    rec = _bodies(n)
    ref = xo.rigid38(rec).astype(np.float32)
    dt = 1 / 60 / 25
    # frame
    out = np.zeros(n, dtype=_abi.FRAME_DTYPE)
    _abi.check(L().xpbd_rigid_frame_batch(_abi.ptr(rec), n, _abi.ptr(out)))
    want = np.zeros((n, 7), dtype=np.float32)
    for i in range(n):
        xo.lib.xo_rigid_frame_f32(C.c_void_p(ref[i].ctypes.data), C.c_void_p(want[i].ctypes.data))
    assert np.array_equal(_f7(out), want.astype(np.float64))
    # integrate
    got = rec.copy()
    _abi.check(L().xpbd_rigid_integrate_batch(_abi.ptr(got), n, dt))
    for i in range(n):
        xo.rigid_op("integrate", ref[i], dt)
    assert np.array_equal(xo.rigid38(got).astype(np.float32), ref)  # constants stay f64 on the host record
    # derive
    past = _f7(_frames(n, 7)).copy()
    _abi.check(L().xpbd_rigid_derive_batch(_abi.ptr(got), _abi.ptr(past), n, dt))
    for i in range(n):
        xo.rigid_op("derive", ref[i], past[i, :3], past[i, 3:], dt)
    assert np.array_equal(xo.rigid38(got).astype(np.float32), ref)  # constants stay f64 on the host record
    # apply_impulse
    rng = np.random.default_rng(3)
    ip = np.concatenate([rng.uniform(-1e-2, 1e-2, (n, 3)), rng.uniform(-1, 1, (n, 3))], axis=1)
    _abi.check(L().xpbd_rigid_apply_impulse_batch(_abi.ptr(got), _abi.ptr(ip), n))
    for i in range(n):
        xo.rigid_op("apply_impulse", ref[i], ip[i, :3], ip[i, 3:])
    assert np.array_equal(xo.rigid38(got).astype(np.float32), ref)  # constants stay f64 on the host record


@pytest.mark.parametrize("hull", ["cube", "icosa", "tetra"])
def test_ground_and_solve(hull):
    _, poly = HULLS[hull]
    n = 300
    rec = _bodies(n, seed=5)
    past = _frames(n, 11)
    past["position"] = rec["position"] + 0.01
    nv = len(poly.vertices)
    cons = np.zeros((n, nv), dtype=_abi.CONSTRAINT_DTYPE)
    counts = np.zeros(n, dtype=np.int32)
    c = poly.as_c()
    _abi.check(L().xpbd_ground_batch(_abi.ptr(rec), _abi.ptr(past), n, C.byref(c), _abi.ptr(cons), _abi.ptr(counts)))
    ref = xo.rigid38(rec).astype(np.float32)
    p7 = _f7(past).astype(np.float32)
    total = 0
    for i in range(n):
        want = xo.ground(ref[i], p7[i], poly.vertices)
        assert counts[i] == len(want)
        got = np.concatenate([cons["contact0"][i, :counts[i]], cons["contact1"][i, :counts[i]], cons["distance"][i, :counts[i], None]], axis=1)
        assert np.array_equal(got, want.astype(np.float64)), i
        total += len(want)
    assert total > n // 4  # the scene does produce contacts
    # solve: sweep those constraints
    dt = 1 / 60 / 25
    got = rec.copy()
    _abi.check(L().xpbd_solve_batch(_abi.ptr(got), n, _abi.ptr(cons), _abi.ptr(counts), nv, dt, 1e-6))
    for i in range(n):
        k = counts[i]
        c7 = np.concatenate([cons["contact0"][i, :k], cons["contact1"][i, :k], cons["distance"][i, :k, None]], axis=1)
        xo.solve(ref[i], c7.astype(np.float32), np.float32(dt))
    assert np.array_equal(xo.rigid38(got).astype(np.float32), ref)  # constants stay f64 on the host record


PAIRS = [("cube", "cube"), ("cube", "icosa"), ("icosa", "tetra"), ("tetra", "cube"), ("icosa", "icosa")]


@pytest.mark.parametrize("a,b", PAIRS)
def test_face_and_edge_queries(a, b):
    (ka, pa), (kb, pb) = HULLS[a], HULLS[b]
    n = 400
    fa, fb = _frames(n, 21), _frames(n, 22)
    # a few exactly aligned / touching configurations: ties and parallel edges (NaN axes)
    fa["rotation"][:8] = (0, 0, 0, 1); fb["rotation"][:8] = (0, 0, 0, 1)
    fa["position"][:8] = 0; fb["position"][:8] = [(0, 0, 0.999), (0, 0, 1.0), (0.5, 0, 0.99), (1, 1, 1), (0.3, 0.3, 0.3), (0, 0, 2), (0, 1.0, 0), (-0.99, 0, 0)]
    dist = np.zeros(n); face = np.zeros(n, dtype=np.int32)
    ca, cb = pa.as_c(), pb.as_c()
    _abi.check(L().xpbd_face_axes_separation_batch(_abi.ptr(fa), _abi.ptr(fb), n, C.byref(ca), C.byref(cb), _abi.ptr(dist), _abi.ptr(face)))
    ha, hb = xo.builtin_hull(ka), xo.builtin_hull(kb)
    f7a, f7b = _f7(fa), _f7(fb)
    for i in range(n):
        d, f = xo.face_axes_separation(f7a[i], f7b[i], ha, hb, np.float32)
        assert (np.float32(dist[i]) == d or (d == 0 and dist[i] == 0)) and face[i] == f, (i, dist[i], d, face[i], f)
    edist = np.zeros(n); edges = np.zeros((n, 2), dtype=np.int32)
    _abi.check(L().xpbd_edge_axes_separation_batch(_abi.ptr(fa), _abi.ptr(fb), n, C.byref(ca), C.byref(cb), _abi.ptr(edist), _abi.ptr(edges)))
    for i in range(n):
        d, e = xo.edge_axes_separation(f7a[i], f7b[i], ha, hb, np.float32)
        assert np.float32(edist[i]) == d and tuple(edges[i]) == e, (i, edist[i], d, edges[i], e)


@pytest.mark.parametrize("offset", [(0.0, 0.0, 0.0), (511.3, -297.1, 88.7)], ids=["origin", "far"])
@pytest.mark.parametrize("a,b", PAIRS)
def test_sat_manifold(a, b, offset):
    """The finished sat through the STEP's edge query (witness cull + reference arithmetic on the survivors): same feature,
    same contacts as the oracle's full scan, near the origin and far from it (the cull's rounding margin scales with
    the coordinates)."""
    (_, pa), (_, pb) = HULLS[a], HULLS[b]
    n = 600
    fa, fb = _frames(n, 31, spread=0.7), _frames(n, 32, spread=0.7)
    # stacked / shallow configurations
    k = n // 3
    rng = np.random.default_rng(9)
    small = scenes.random_unit_quaternions(rng, k); small[:, :3] *= 0.02; small /= np.linalg.norm(small, axis=1)[:, None]
    fa["rotation"][:k] = (0, 0, 0, 1); fa["position"][:k] = 0
    fb["rotation"][:k] = small; fb["position"][:k] = np.stack([rng.uniform(-0.3, 0.3, k), rng.uniform(-0.3, 0.3, k), rng.uniform(0.9, 1.02, k)], axis=1)
    fa["position"] += np.array(offset); fb["position"] += np.array(offset)
    cons = np.zeros((n, 4), dtype=_abi.CONSTRAINT_DTYPE); counts = np.zeros(n, dtype=np.int32); feat = np.zeros(n, dtype=np.int32)
    ca, cb = pa.as_c(), pb.as_c()
    _abi.check(L().xpbd_sat_manifold_batch(_abi.ptr(fa), _abi.ptr(fb), n, C.byref(ca), C.byref(cb), _abi.ptr(cons), _abi.ptr(counts), _abi.ptr(feat)))
    f7a, f7b = _f7(fa), _f7(fb)
    seen = set()
    for i in range(n):
        pts, tg, feature, inc = xw.sat_manifold(f7a[i], f7b[i], pa, pb, np.float32)
        seen.add(feature)
        assert feat[i] == feature and counts[i] == len(pts), (i, feat[i], feature, counts[i], len(pts))
        if len(pts):
            assert np.array_equal(cons["contact0"][i, :len(pts)], pts.astype(np.float64)), i
            assert np.array_equal(cons["contact1"][i, :len(pts)], tg.astype(np.float64)), i
            assert (cons["rigid"][i, :len(pts)] == inc).all()
    assert {0, 1}.issubset(seen) and len(seen) >= 3, seen  # separated, face contacts and at least one more feature


def test_fp32_peak_microbench_runs():
    t = C.c_double(0)
    _abi.check(L().xpbd_fp32_peak_microbench(0, 0, C.byref(t)))
    assert 10.0 < t.value < 100.0  # ~37 TFLOP/s nominal without FMA
