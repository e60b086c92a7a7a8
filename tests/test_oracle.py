"""CPU oracle pinned against the known answers SURVEY.md §8 row C lists (the reference has no tests or
golden vectors of its own: PARITY UNPINNED) and against the committed regression fixtures."""
import json
import os

import numpy as np
import pytest

import oracle_lib as xo

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
TETRA, CUBE, ICOSA = 0, 1, 2


def test_cube_mass_properties():
    m = xo.rigid_metrics(xo.builtin_hull(CUBE), 0.1)
    assert m[0] == pytest.approx(0.1, abs=1e-15) and m[1] == pytest.approx(1.0, abs=1e-15)
    np.testing.assert_allclose(m[2:5], [0.5, 0.5, 0.5], atol=1e-15)
    np.testing.assert_allclose(m[5:].reshape(3, 3), np.eye(3) / 60.0, atol=1e-15)


def test_scaled_tetrahedron_mass_properties():
    v, e, fo, fi, c = xo.builtin_hull(TETRA)
    m = xo.rigid_metrics((0.5 * v, e, fo, fi, 0.5 * c), 5.0)
    assert m[0] == pytest.approx(0.10416666666666667, rel=1e-13)
    assert m[1] == pytest.approx(0.020833333333333332, rel=1e-13)
    np.testing.assert_allclose(m[2:5], [0.125] * 3, atol=1e-15)
    J = m[5:].reshape(3, 3)
    np.testing.assert_allclose(np.diag(J), [1.953125e-3] * 3, rtol=1e-12)
    np.testing.assert_allclose(J[0, 1], 3.2552083333e-4, rtol=1e-9)
    rc, r = xo.rigid_new(m)
    assert rc == 0
    Jinv = r[1:10].reshape(3, 3)
    np.testing.assert_allclose(np.diag(Jinv), [537.6] * 3, rtol=1e-12)
    np.testing.assert_allclose(Jinv[0, 1], -76.8, rtol=1e-12)


def test_icosahedron_mass_properties():
    m = xo.rigid_metrics(xo.builtin_hull(ICOSA), 1.0)
    assert m[0] == pytest.approx(2.5361507101, rel=1e-10)
    np.testing.assert_allclose(m[2:5], 0.0, atol=1e-15)
    np.testing.assert_allclose(np.diag(m[5:].reshape(3, 3)), [0.7340703576] * 3, rtol=1e-9)


def test_cube_planes_in_face_order():
    pl = xo.hull_planes(xo.builtin_hull(CUBE))
    expect = [(0, 0, -1, 0), (0, 0, 1, 1), (0, -1, 0, 0), (1, 0, 0, 1), (0, 1, 0, 1), (-1, 0, 0, 0)]
    np.testing.assert_allclose(pl, expect, atol=0)


def test_singular_inertia_reports_like_the_reference_panics():
    m = np.zeros(14); m[0] = 1.0
    rc, _ = xo.rigid_new(m)
    assert rc == -1


def test_world_new_matches_survey():
    w = xo.world_new()
    a, b = w
    assert a[0] == pytest.approx(10.0) and np.allclose(a[1:10].reshape(3, 3), 60 * np.eye(3))
    np.testing.assert_allclose(a[31:34], [0, 0, 4]); np.testing.assert_allclose(a[22:25], [0, 2.5, 0])
    np.testing.assert_allclose(a[25:28], [-4, 1, 0]); np.testing.assert_allclose(a[10:13], [0, 0, -2])
    # b: Euler(10, 15, 5 deg) -> (s, x, y, z) ~ (0.986236, 0.092000, 0.126137, 0.054447)
    np.testing.assert_allclose(b[[37, 34, 35, 36]], [0.986236, 0.092000, 0.126137, 0.054447], atol=1e-6)


def test_free_fall_closed_form():
    """No contact: z = 4 - 20*dt_s^2*n(n+1)/2, y = 2.5*n*dt_s."""
    w = xo.world_new()
    xo.world_integrate(w, 1 / 60, 1)
    dts = 1 / 60 / 25
    assert w[0, 33] == pytest.approx(4 - 20 * dts * dts * 25 * 26 / 2, abs=1e-12)
    assert w[0, 32] == pytest.approx(2.5 * 25 * dts, abs=1e-12)
    assert w[0, 33] == pytest.approx(3.9971111, abs=1e-7) and w[0, 32] == pytest.approx(0.0416667, abs=1e-7)


def test_survey_f64_checkpoints_body_a():
    """Checkpoints produced by the survey's independent throwaway restatement (NOT by the reference binary).
    Only frames <= 60 are usable: the tumble is chaotic (a 1e-15 perturbation reaches 1e-2 by frame 80,
    see DESIGN.md), so the survey's frame-100/600 values are not reproducible by any other rounding order."""
    w = xo.world_new()
    got = {}
    for f in range(1, 61):
        xo.world_integrate(w, 1 / 60, 1)
        got[f] = (w[0, 31:34].copy(), w[0, [37, 34, 35, 36]].copy())
    np.testing.assert_allclose(got[10][0], [0, 0.416667, 3.721111], atol=1e-6)
    np.testing.assert_allclose(got[10][1], [0.941564, -0.326776, 0.081694, 0], atol=1e-6)
    np.testing.assert_allclose(got[30][0], [0, 1.25, 1.496667], atol=1e-6)
    np.testing.assert_allclose(got[30][1], [0.514466, -0.831908, 0.207977, 0], atol=1e-6)
    np.testing.assert_allclose(got[60][0], [-0.428683, 3.051182, 0.268415], atol=1e-6)


def test_survey_f64_late_checkpoints_same_outcome_within_chaos():
    """SURVEY.md section 8(c)(3), frames 100 / 600 and the 29 559-constraint count.  These are past the chaotic tumble
    (frames 60-90: a 1e-15 perturbation grows to centimetres), so two f64 implementations with different rounding
    order agree only in OUTCOME, not in digits: the cube must settle on the same face (orientation within 0.05 rad of
    the survey's), within 5 cm of its resting place, on the ground, having applied the same number of constraints to 3 %.
    This is a sanity anchor, not a pin — the digits are pinned by test_known_answer.py and the frames <= 60 above."""
    import parity
    a = xo.world_new()[0:1].copy()
    cube = xo.builtin_hull(CUBE)[0]
    n = xo.step_bodies(a, cube, 1 / 60, 25, 100, threads=1)
    q100 = np.array([[0.054340, 0.705020, 0.054339, -0.705012]])  # survey: (s, x, y, z) = (-0.705012, 0.054340, 0.705020, 0.054339)
    err = min(parity.rot_error(a[:, 34:38], q100)[0], parity.rot_error(-a[:, 34:38], q100)[0])
    assert err < 0.05, err
    np.testing.assert_allclose(a[0, 31:34], [-0.763723, 3.856061, -4.6e-7], atol=0.05)
    n += xo.step_bodies(a, cube, 1 / 60, 25, 500, threads=1)
    np.testing.assert_allclose(a[0, 31:34], [-0.763889, 3.856085, 3.3e-6], atol=0.05)
    assert abs(a[0, 33]) < 1e-4
    assert abs(n - 29_559) < 0.03 * 29_559, n


def test_cube_comes_to_rest_on_a_face():
    w = xo.world_new()
    xo.world_integrate(w, 1 / 60, 600)
    a = w[0]
    assert abs(a[33]) < 1e-4  # object-space origin (a cube corner) ends on the ground plane: a face is down
    assert np.linalg.norm(a[22:25]) < 1e-2 and np.linalg.norm(a[25:28]) < 5e-2  # residual creep only
    assert np.isfinite(w).all()


def test_f32_and_f64_instantiations_agree_early():
    w64 = xo.world_new()
    w32 = w64.astype(np.float32)
    xo.world_integrate(w64, 1 / 60, 5); xo.world_integrate(w32, 1 / 60, 5)
    np.testing.assert_allclose(w32[:, 31:38], w64[:, 31:38], atol=2e-3)  # 2e-3 m apart by frame 10


def test_threaded_step_equals_serial():
    rng = np.random.default_rng(0)
    base = xo.world_new()[0].astype(np.float32)
    r = np.tile(base, (64, 1)); r[:, 33] += rng.random(64).astype(np.float32)
    verts = xo.builtin_hull(CUBE)[0]
    r1, r4 = r.copy(), r.copy()
    n1 = xo.step_bodies(r1, verts, 1 / 60, 25, 40, threads=1)
    n4 = xo.step_bodies(r4, verts, 1 / 60, 25, 40, threads=4)
    assert n1 == n4 and np.array_equal(r1, r4)


@pytest.mark.parametrize("name", ["c1_f32", "c1_f64"])
def test_golden_regression(name):
    """Fixtures generated by tests/golden/make_golden.py from this oracle (regression anchors: they pin the
    oracle against accidental edits, not against the reference, which cannot be run here)."""
    with open(os.path.join(GOLDEN, name + ".json")) as f:
        g = json.load(f)
    dtype = np.float32 if name.endswith("f32") else np.float64
    w = xo.world_new().astype(dtype)
    frame = 0
    for cp in g["checkpoints"]:
        xo.world_integrate(w, dtype(1 / 60) if dtype == np.float32 else 1 / 60, cp["frame"] - frame)
        frame = cp["frame"]
        for body, key in ((0, "a"), (1, "b")):
            got = w[body].astype(np.float64)
            want = np.array([float.fromhex(h) for h in cp[key]])
            assert np.array_equal(got, want), f"{name} frame {frame} body {key}"
