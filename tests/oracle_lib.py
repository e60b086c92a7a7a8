"""ctypes access to the CPU oracle (oracle/_build/liboracle.so) — TEST INFRASTRUCTURE ONLY."""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
_L = C.CDLL(os.path.join(ROOT, "oracle", "_build", "liboracle.so"))
_VP = C.c_void_p


def _p(a):
    return None if a is None else C.c_void_p(a.ctypes.data)


for name in ("xo_step_bodies_f32", "xo_step_bodies_f64", "xo_world_integrate_f64", "xo_world_integrate_f32"):
    getattr(_L, name).restype = C.c_uint64
_L.xo_step_bodies_f32.argtypes = [_VP, C.c_int64, _VP, C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
_L.xo_step_bodies_f64.argtypes = [_VP, C.c_int64, _VP, C.c_int, C.c_double, C.c_int, C.c_int, C.c_int]
_L.xo_world_integrate_f64.argtypes = [_VP, C.c_double, C.c_int]
_L.xo_world_integrate_f32.argtypes = [_VP, C.c_float, C.c_int]
_L.xo_face_axes_separation_f32.restype = C.c_float
_L.xo_face_axes_separation_f64.restype = C.c_double
_L.xo_edge_axes_separation_f32.restype = C.c_float
_L.xo_edge_axes_separation_f64.restype = C.c_double

# C layouts of the oracle's own entry points (oracle/xo_world_capi.inc: XoPolytope, XoJoint; Rigid = 38 scalars).  They
# equal the product's records by construction (tests/test_abi.py checks it), but the oracle side declares them itself so
# that nothing here imports or maps the product library.
RIGID_DTYPE = np.dtype([("inverse_mass", "f8"), ("inverse_inertia", "f8", (9,)), ("external_force", "f8", (3,)),
                        ("internal_force", "f8", (3,)), ("external_torque", "f8", (3,)), ("internal_torque", "f8", (3,)),
                        ("velocity", "f8", (3,)), ("angular_velocity", "f8", (3,)), ("center_of_mass", "f8", (3,)),
                        ("position", "f8", (3,)), ("rotation", "f8", (4,))])
JOINT_DTYPE = np.dtype([("body_a", "i4"), ("body_b", "i4"), ("anchor_a", "f8", (3,)), ("anchor_b", "f8", (3,)), ("distance", "f8")])


class XoPolytope(C.Structure):
    _fields_ = [("vertices", C.POINTER(C.c_double)), ("n_vertices", C.c_int32), ("edges", C.POINTER(C.c_int32)),
                ("n_edges", C.c_int32), ("face_offsets", C.POINTER(C.c_int32)), ("face_indices", C.POINTER(C.c_int32)),
                ("n_faces", C.c_int32), ("centroid", C.c_double * 3)]


class Hull:
    """Plain holder of a Polytope's arrays (src/geometry.rs:81-93) for the oracle's entry points.  Anything with
    .vertices / .edges / .faces / .centroid (the product's Polytope included) converts through Hull.of()."""

    def __init__(self, vertices, edges, faces, centroid):
        self.vertices = np.ascontiguousarray(vertices, dtype=np.float64).reshape(-1, 3)
        self.edges = np.ascontiguousarray(edges, dtype=np.int32).reshape(-1, 2)
        self.faces = [list(map(int, f)) for f in faces]
        self.centroid = np.ascontiguousarray(centroid, dtype=np.float64).reshape(3)
        self.face_offsets = np.zeros(len(self.faces) + 1, dtype=np.int32)
        self.face_offsets[1:] = np.cumsum([len(f) for f in self.faces])
        self.face_indices = np.ascontiguousarray([v for f in self.faces for v in f], dtype=np.int32)

    @staticmethod
    def of(h) -> "Hull":
        return h if isinstance(h, Hull) else Hull(h.vertices, h.edges, h.faces, h.centroid)

    @staticmethod
    def builtin(kind: int) -> "Hull":
        v, e, fo, fi, c = builtin_hull(kind)
        return Hull(v, e, [fi[fo[i]:fo[i + 1]].tolist() for i in range(len(fo) - 1)], c)

    def scaled(self, k: float) -> "Hull":  # `f64 * Polytope`, src/geometry.rs:296-307
        return Hull(k * self.vertices, self.edges, self.faces, k * self.centroid)

    def tuple5(self):
        return self.vertices, self.edges, self.face_offsets, self.face_indices, self.centroid

    def as_c(self) -> XoPolytope:
        p = XoPolytope()
        p.vertices = self.vertices.ctypes.data_as(C.POINTER(C.c_double)); p.n_vertices = len(self.vertices)
        p.edges = self.edges.ctypes.data_as(C.POINTER(C.c_int32)); p.n_edges = len(self.edges)
        p.face_offsets = self.face_offsets.ctypes.data_as(C.POINTER(C.c_int32))
        p.face_indices = self.face_indices.ctypes.data_as(C.POINTER(C.c_int32))
        p.n_faces = len(self.faces)
        p.centroid[:] = self.centroid.tolist()
        return p


RIGID_FIELDS = [("inverse_mass", 0, 1), ("inverse_inertia", 1, 9), ("external_force", 10, 3), ("internal_force", 13, 3),
                ("external_torque", 16, 3), ("internal_torque", 19, 3), ("velocity", 22, 3), ("angular_velocity", 25, 3),
                ("center_of_mass", 28, 3), ("position", 31, 3), ("rotation", 34, 4)]


def hardware_threads() -> int:
    return int(_L.xo_hardware_threads())


def rigid38(rec: np.ndarray) -> np.ndarray:
    """xpbd_rigid records (structured f64) -> plain (n, 38) f64 view-copy."""
    return np.ascontiguousarray(rec).view(np.float64).reshape(-1, 38).copy()


def step_bodies(rigids: np.ndarray, verts: np.ndarray, dt: float, substeps: int, frames: int, threads: int = 1) -> int:
    """solver::step on n bodies in place. dtype of `rigids` (f32/f64, shape (n,38)) selects the instantiation."""
    assert rigids.flags["C_CONTIGUOUS"] and rigids.shape[-1] == 38
    v = np.ascontiguousarray(verts, dtype=rigids.dtype).reshape(-1, 3)
    n = rigids.size // 38
    if rigids.dtype == np.float32:
        return int(_L.xo_step_bodies_f32(_p(rigids), n, _p(v), len(v), np.float32(dt), substeps, frames, threads))
    return int(_L.xo_step_bodies_f64(_p(rigids), n, _p(v), len(v), float(dt), substeps, frames, threads))


def world_new() -> np.ndarray:
    w = np.zeros((2, 38))
    _L.xo_world_new_f64(_p(w))
    return w


def world_integrate(ab: np.ndarray, dt: float, frames: int) -> int:
    if ab.dtype == np.float32:
        return int(_L.xo_world_integrate_f32(_p(ab), np.float32(dt), frames))
    return int(_L.xo_world_integrate_f64(_p(ab), float(dt), frames))


def builtin_hull(kind: int):
    n = np.zeros(4, dtype=np.int32)
    _L.xo_builtin_hull_f64(kind, None, None, None, None, None, _p(n))
    verts = np.zeros((n[0], 3)); edges = np.zeros((n[1], 2), dtype=np.int32)
    foff = np.zeros(n[2] + 1, dtype=np.int32); fidx = np.zeros(n[3], dtype=np.int32); cen = np.zeros(3)
    _L.xo_builtin_hull_f64(kind, _p(verts), _p(edges), _p(foff), _p(fidx), _p(cen), _p(n))
    return verts, edges, foff, fidx, cen


def _hull_args(h, dtype):
    verts, edges, foff, fidx, cen = h
    v = np.ascontiguousarray(verts, dtype=dtype); c = np.ascontiguousarray(cen, dtype=dtype)
    return (v, c), (_p(v), len(v), _p(edges), len(edges), _p(foff), _p(fidx), len(foff) - 1, _p(c))


def rigid_metrics(h, density: float) -> np.ndarray:
    keep, args = _hull_args(h, np.float64)
    out = np.zeros(14)
    rc = _L.xo_rigid_metrics_f64(*args, C.c_double(density), _p(out))
    assert rc == 0
    return out


def hull_planes(h, dtype=np.float64) -> np.ndarray:
    keep, args = _hull_args(h, dtype)
    out = np.zeros((len(h[2]) - 1, 4), dtype=dtype)
    fn = _L.xo_hull_planes_f32 if dtype == np.float32 else _L.xo_hull_planes_f64
    assert fn(*args, _p(out)) == 0
    return out


def rigid_new(metrics14: np.ndarray):
    out = np.zeros(38)
    rc = _L.xo_rigid_new_f64(_p(np.ascontiguousarray(metrics14)), _p(out))
    return rc, out


def rigid_op(name: str, r: np.ndarray, *args):
    """integrate / derive / apply_impulse / frame on one rigid38 row (dtype selects f32/f64)."""
    suf = "f32" if r.dtype == np.float32 else "f64"
    fn = getattr(_L, f"xo_rigid_{name}_{suf}")
    sc = C.c_float if r.dtype == np.float32 else C.c_double
    cargs = []
    for a in args:
        if isinstance(a, np.ndarray):
            a = np.ascontiguousarray(a, dtype=r.dtype); cargs.append((a, _p(a)))
        else:
            cargs.append((None, sc(a)))
    fn(_p(r), *[c[1] for c in cargs])
    return [c[0] for c in cargs]


def ground(r: np.ndarray, past7: np.ndarray, verts: np.ndarray) -> np.ndarray:
    v = np.ascontiguousarray(verts, dtype=r.dtype).reshape(-1, 3)
    out = np.zeros((len(v), 7), dtype=r.dtype)
    fn = _L.xo_ground_f32 if r.dtype == np.float32 else _L.xo_ground_f64
    n = fn(_p(r), _p(np.ascontiguousarray(past7, dtype=r.dtype)), _p(v), len(v), _p(out))
    return out[:n]


def solve(r: np.ndarray, cons7: np.ndarray, dt: float) -> None:
    c = np.ascontiguousarray(cons7, dtype=r.dtype).reshape(-1, 7)
    if r.dtype == np.float32:
        _L.xo_solve_f32(_p(r), _p(c), len(c), C.c_float(dt))
    else:
        _L.xo_solve_f64(_p(r), _p(c), len(c), C.c_double(dt))


def face_axes_separation(fa, fb, hull_a, hull_b, dtype=np.float32):
    keep, args = _hull_args(hull_a, dtype)
    vb = np.ascontiguousarray(hull_b[0], dtype=dtype)
    face = C.c_int32(-1)
    fa = np.ascontiguousarray(fa, dtype=dtype); fb = np.ascontiguousarray(fb, dtype=dtype)
    fn = _L.xo_face_axes_separation_f32 if dtype == np.float32 else _L.xo_face_axes_separation_f64
    d = fn(_p(fa), _p(fb), *args, _p(vb), len(vb), C.byref(face))
    return d, face.value


def edge_axes_separation(fa, fb, hull_a, hull_b, dtype=np.float32):
    va = np.ascontiguousarray(hull_a[0], dtype=dtype); ca = np.ascontiguousarray(hull_a[4], dtype=dtype)
    vb = np.ascontiguousarray(hull_b[0], dtype=dtype)
    ea = np.ascontiguousarray(hull_a[1], dtype=np.int32); eb = np.ascontiguousarray(hull_b[1], dtype=np.int32)
    pair = np.zeros(2, dtype=np.int32)
    fa = np.ascontiguousarray(fa, dtype=dtype); fb = np.ascontiguousarray(fb, dtype=dtype)
    fn = _L.xo_edge_axes_separation_f32 if dtype == np.float32 else _L.xo_edge_axes_separation_f64
    d = fn(_p(fa), _p(fb), _p(va), len(va), _p(ea), len(ea), _p(ca), _p(vb), len(vb), _p(eb), len(eb), _p(pair))
    return d, (int(pair[0]), int(pair[1]))


lib = _L
