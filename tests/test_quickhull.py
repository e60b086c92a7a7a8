"""xpbd_quickhull (README.md:20 of the reference): host-side f64, no GPU needed."""
import numpy as np
import pytest

import constraint_solver_b200 as cs
from constraint_solver_b200 import _abi, scenes


def _check_hull(h, pts):
    V, E, F = len(h.vertices), len(h.edges), len(h.faces)
    assert V - E + F == 2, (V, E, F)  # Euler
    planes = h.planes()  # outward from the centroid (src/geometry.rs:262-271)
    scale = np.ptp(pts, axis=0).max()
    for pl in planes:
        d = pts @ pl.normal - pl.displacement
        assert d.max() <= 1e-7 * scale, d.max()  # every input point inside every face plane
    for f, pl in zip(h.faces, planes):  # face vertices coplanar
        assert np.abs(h.vertices[f] @ pl.normal - pl.displacement).max() <= 1e-7 * scale
    # every edge is shared by exactly two faces (manifold, src/geometry.rs:79)
    count = {}
    for f in h.faces:
        for a, b in zip(f, f[1:] + f[:1]):
            count[(min(a, b), max(a, b))] = count.get((min(a, b), max(a, b)), 0) + 1
    assert set(count) == {(min(a, b), max(a, b)) for a, b in h.edges.tolist()}
    assert all(v == 2 for v in count.values())
    assert np.allclose(h.centroid, h.vertices.mean(axis=0))


def test_cube_with_interior_and_duplicate_points_gives_quads():
    cube = cs.Polytope.new_cube()
    pts = np.vstack([cube.vertices, [[0.5, 0.5, 0.5], [0.2, 0.3, 0.4], [0.5, 0.5, 0.0], [1.0, 1.0, 1.0]]])
    h = cs.Polytope.from_points(pts)
    assert (len(h.vertices), len(h.edges), len(h.faces)) == (8, 12, 6)
    assert all(len(f) == 4 for f in h.faces)
    _check_hull(h, pts)
    m = h.rigid_metrics(0.1)
    assert m.mass == pytest.approx(0.1, rel=1e-12) and np.allclose(m.center_of_mass, 0.5)


def test_tetrahedron_and_volume_against_scipy():
    from scipy.spatial import ConvexHull
    h = cs.Polytope.from_points(cs.Polytope.new_tetrahedron().vertices)
    assert (len(h.vertices), len(h.edges), len(h.faces)) == (4, 6, 4)
    rng = np.random.default_rng(7)
    for n in (8, 20, 64, 200):
        pts = rng.normal(size=(n, 3))
        h = cs.Polytope.from_points(pts)
        _check_hull(h, pts)
        assert h.rigid_metrics(1.0).volume == pytest.approx(ConvexHull(pts).volume, rel=1e-9)
        assert len(h.vertices) == len(ConvexHull(pts).vertices)


def test_degenerate_inputs_are_errors_not_crashes():
    for pts in ([[0, 0, 0]] * 5, [[0, 0, 0], [1, 0, 0], [2, 0, 0], [3, 0, 0]], [[0, 0, 0], [1, 0, 0], [0, 1, 0], [1, 1, 0]],
                [[0, 0, 0], [1, 0, 0], [0, 1, 0]]):
        with pytest.raises(cs.XpbdError) as e:
            cs.Polytope.from_points(np.array(pts, dtype=float))
        assert e.value.status == _abi.XPBD_ERR_INVALID
    with pytest.raises(cs.XpbdError):
        cs.Polytope.from_points(np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [np.nan, 0, 1]]))


def test_c4_hull_table_is_valid_for_the_multi_body_path():
    hulls = scenes.c4_hulls()
    assert len(hulls) == 8
    for h in hulls:
        assert len(h.vertices) <= 64 and max(len(f) for f in h.faces) <= 8 and len(h.edges) > 0
        _check_hull(h, h.vertices)
        assert h.rigid_metrics(0.3).mass > 0
