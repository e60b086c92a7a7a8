"""The oracle-side scene generators (tests/oracle_scenes.py, what bench.py's CPU arms step without importing the
product) produce the same worlds, bit for bit, as the product-side generators the GPU arm uploads."""
import numpy as np

import oracle_lib as xo
import oracle_scenes as xs
from constraint_solver_b200 import _abi, scenes


def test_record_layouts_declared_on_both_sides_agree():
    assert xo.RIGID_DTYPE == _abi.RIGID_DTYPE and xo.JOINT_DTYPE == _abi.JOINT_DTYPE
    import ctypes as C
    assert C.sizeof(xo.XoPolytope) == C.sizeof(_abi.PolytopeC)
    assert [f[0] for f in xo.XoPolytope._fields_] == [f[0] for f in _abi.PolytopeC._fields_]


def test_c2_and_c3_scenes_equal_the_product_side_generators():
    for first in (0, 777):
        assert np.array_equal(xs.c2_perturbed(300, first).reshape(-1, 38), xo.rigid38(scenes.c2_perturbed(300, first)))
        assert np.array_equal(xs.c3_stack(200, first).reshape(-1, 38), xo.rigid38(scenes.c3_stack(200, first)))


def test_c4_fixture_and_scene_equal_the_product_side_generators():
    fresh, fixture = scenes.c4_hulls(), xs.c4_hulls()
    assert len(fresh) == len(fixture) == 8
    for a, b in zip(fresh, fixture):
        assert np.array_equal(a.vertices, b.vertices) and np.array_equal(a.edges, b.edges)
        assert a.faces == b.faces and np.array_equal(a.centroid, b.centroid)
    init, hulls, ids, joints = scenes.c4_pile(9, first_world=5)
    oinit, ohulls, oids, ojoints = xs.c4_pile(9, first_world=5)
    assert np.array_equal(oinit.reshape(-1, 38), xo.rigid38(init))
    assert np.array_equal(ids, oids) and joints.tobytes() == ojoints.tobytes()
