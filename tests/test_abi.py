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


def _header_prototypes():
    """name -> number of parameters, from include/xpbd_b200.h."""
    text = open(os.path.join(ROOT, "include", "xpbd_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    out = {}
    for m in re.finditer(r"\b(xpbd_[a-z0-9_]+)\s*\(([^;{}]*?)\)\s*;", text, flags=re.S):
        args = m.group(2).strip()
        out[m.group(1)] = 0 if args in ("", "void") else args.count(",") + 1
    return out


def test_rust_ffi_declares_every_header_symbol_with_the_same_arity():
    """rust/src/ffi.rs (the shim a maintainer drops into the crate; it cannot be compiled here — no cargo/rustc) must
    bind exactly the header's entry points, each with the header's number of parameters, and mirror the record sizes."""
    text = open(os.path.join(ROOT, "rust", "src", "ffi.rs")).read()
    block = text[text.index('extern "C" {'):]
    block = block[:block.index("\n}\n")]
    rust = {}
    for m in re.finditer(r"pub fn (xpbd_[a-z0-9_]+)\s*\((.*?)\)\s*->", block, flags=re.S):
        args = m.group(2).strip()
        rust[m.group(1)] = 0 if not args else len([a for a in args.split(",") if a.strip()])
    header = _header_prototypes()
    assert set(rust) == set(header), set(rust) ^ set(header)
    assert rust == header, {k: (rust[k], header[k]) for k in header if rust[k] != header[k]}
    # every Python binding has the same arity too
    assert {k: len(v[1]) for k, v in _abi._PROTOTYPES.items()} == header
    # #[repr(C)] records: same field lists as the numpy / ctypes mirrors (names in declaration order)
    def rust_fields(name):
        body = text[text.index(f"pub struct {name} {{"):]
        body = body[:body.index("}")]
        return re.findall(r"pub (\w+):", body)
    assert rust_fields("xpbd_rigid") == list(_abi.RIGID_DTYPE.names)
    assert rust_fields("xpbd_frame") == list(_abi.FRAME_DTYPE.names)
    assert rust_fields("xpbd_constraint") == list(_abi.CONSTRAINT_DTYPE.names)
    assert rust_fields("xpbd_joint") == list(_abi.JOINT_DTYPE.names)
    assert rust_fields("xpbd_rigid_metrics") == list(_abi.METRICS_DTYPE.names)
    assert rust_fields("xpbd_step_stats") == [f for f, _ in _abi.StepStatsC._fields_]
    assert rust_fields("xpbd_batch_desc") == [f for f, _ in _abi.BatchDescC._fields_]
    assert rust_fields("xpbd_polytope") == [f for f, _ in _abi.PolytopeC._fields_]


def test_rust_shim_covers_the_reference_module_surface():
    """Names a maintainer's code calls exist in the shim sources."""
    src = {f: open(os.path.join(ROOT, "rust", "src", f)).read() for f in os.listdir(os.path.join(ROOT, "rust", "src"))}
    for f, names in {"rigid.rs": ["pub fn new(", "pub fn frame(", "pub fn integrate(", "pub fn derive(", "pub fn apply_impulse("],
                     "solver.rs": ["pub fn step(", "pub fn solve("], "world.rs": ["pub fn new(", "pub fn integrate("],
                     "collision.rs": ["pub fn ground(", "pub fn face_axes_separation(", "pub fn edge_axes_separation(", "pub fn sat("],
                     "worlds.rs": ["pub fn set_hulls(", "pub fn set_joints(", "pub fn set_sub_batches(", "pub fn integrate_host(",
                                   "pub fn download_dynamic(", "pub fn upload_dynamic(", "pub fn integrate("]}.items():
        for n in names:
            assert n in src[f], (f, n)


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
