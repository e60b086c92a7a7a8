"""The C-ABI library loads and exports every symbol include/xpbd_b200.h declares; host-side (no GPU)
entry points agree with the oracle; compute entry points fail loudly without a device."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import oracle_lib as xo
import constraint_solver_b200 as cs
from constraint_solver_b200 import _abi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "xpbd_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(xpbd_[a-z0-9_]+)\s*\(", text)))


def test_every_declared_symbol_is_exported_and_bound():
    lib = _abi.load()
    names = _declared_symbols()
    assert len(names) >= 30
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/xpbd_b200.h but not exported"
    assert set(names) == set(_abi.EXPORTED_SYMBOLS), set(names) ^ set(_abi.EXPORTED_SYMBOLS)


def test_record_layouts_match_header():
    assert _abi.RIGID_DTYPE.itemsize == 304 and _abi.FRAME_DTYPE.itemsize == 56
    assert _abi.CONSTRAINT_DTYPE.itemsize == 64 and _abi.JOINT_DTYPE.itemsize == 64
    assert C.sizeof(_abi.StepStatsC) == 64 and C.sizeof(_abi.BatchDescC) == 48


def test_version_string():
    assert b"sm_100a" in _abi.load().xpbd_version()


@pytest.mark.parametrize("kind,poly", [(0, cs.Polytope.new_tetrahedron), (1, cs.Polytope.new_cube), (2, cs.Polytope.new_icosahedron)])
def test_builtin_hulls_and_host_math_match_oracle(kind, poly):
    p = poly()
    v, e, fo, fi, c = xo.builtin_hull(kind)
    assert np.array_equal(p.vertices, v) and np.array_equal(p.edges, e) and np.array_equal(p.centroid, c)
    assert [list(fi[fo[i]:fo[i + 1]]) for i in range(len(fo) - 1)] == p.faces
    m = p.rigid_metrics(0.7)
    mo = xo.rigid_metrics((v, e, fo, fi, c), 0.7)
    got = np.concatenate([[m.mass, m.volume], m.center_of_mass, m.inertia_tensor.reshape(9)])
    assert np.array_equal(got, mo), "product host math (f64) must reproduce the oracle bit for bit"
    planes = np.array([[*pl.normal, pl.displacement] for pl in p.planes()])
    assert np.array_equal(planes, xo.hull_planes((v, e, fo, fi, c)))


def test_world_new_matches_oracle_bitwise():
    from constraint_solver_b200 import scenes
    w = scenes.default_world()
    wo = xo.world_new()
    assert np.array_equal(xo.rigid38(w.a.rec)[0], wo[0])
    np.testing.assert_allclose(xo.rigid38(w.b.rec)[0], wo[1], rtol=0, atol=1e-15)  # libm sin/cos via python vs C


def test_singular_inertia_is_an_error_not_a_panic():
    m = cs.RigidMetrics(1.0, 1.0, np.zeros(3), np.zeros((3, 3)))
    with pytest.raises(cs.XpbdError) as ei:
        cs.Rigid(m)
    assert ei.value.status == _abi.XPBD_ERR_SINGULAR


def test_malformed_hull_rejected():
    bad = cs.Polytope([(0, 0, 0), (1, 0, 0), (0, 1, 0), (0, 0, 1)], [(0, 1)], [[0, 1], [0, 1, 2], [1, 2, 3], [0, 2, 3]], (0.25,) * 3)
    with pytest.raises(cs.XpbdError) as ei:
        bad.rigid_metrics(1.0)
    assert ei.value.status == _abi.XPBD_ERR_INVALID and "fewer than 3" in str(ei.value)


def test_no_cpu_fallback_without_device():
    n = C.c_int32(0)
    rc = _abi.load().xpbd_device_count(C.byref(n))
    if rc == 0 and n.value > 0:
        pytest.skip("a CUDA device is present")
    with pytest.raises(cs.XpbdError) as ei:
        cs.Worlds(4)
    assert ei.value.status == _abi.XPBD_ERR_CUDA and "no CPU fallback" in str(ei.value)


def test_product_does_not_reference_oracle():
    """The product path must not import, link or execute anything under oracle/."""
    pkg = os.path.join(ROOT, "constraint_solver_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".hpp", ".h", ".cpp")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                for needle in ("oracle/", "liboracle", "import oracle", "oracle_lib", "oracle_world", "xo_", "#include \"xo"):
                    assert needle not in text, f"{f} references the oracle ({needle})"
