"""ctypes access to the multi-body part of the CPU oracle (oracle/xo_world.hpp) — TEST INFRASTRUCTURE ONLY."""
import ctypes as C

import numpy as np

import oracle_lib as xo

_L = xo.lib
_VP = C.c_void_p
for _n in ("xo_step_worlds_f32", "xo_step_worlds_f64"):
    getattr(_L, _n).restype = C.c_uint64
_L.xo_step_worlds_f32.argtypes = [_VP, C.c_int64, C.c_int, _VP, C.c_int, _VP, _VP, C.c_int, C.c_double, C.c_double,
                                  C.c_int, C.c_float, C.c_int, C.c_int, C.c_int, _VP]
_L.xo_step_worlds_f64.argtypes = [_VP, C.c_int64, C.c_int, _VP, C.c_int, _VP, _VP, C.c_int, C.c_double, C.c_double,
                                  C.c_int, C.c_double, C.c_int, C.c_int, C.c_int, _VP]


def _hull_array(hulls):
    """hulls: list of oracle_lib.Hull, or of anything with .vertices / .edges / .faces / .centroid (used as data holders)."""
    keep = [xo.Hull.of(h) for h in hulls]
    arr = (xo.XoPolytope * len(keep))()
    for i, h in enumerate(keep):
        arr[i] = h.as_c()
    return arr, keep


def step_worlds(rigids38, B, hulls, dt, substeps, frames, threads=1, hull_ids=None, joints=None, compliance=1e-6,
                friction=1.0, body_body=True):
    """step_world on n_worlds worlds of B bodies in place; `hulls` = list of hulls (oracle_lib.Hull or Polytope-like)."""
    if not isinstance(hulls, (list, tuple)):
        hulls = [xo.Hull.builtin(1)]
    assert rigids38.flags["C_CONTIGUOUS"]
    n_worlds = rigids38.size // (38 * B)
    harr, _keep = _hull_array(hulls)
    ids = None if hull_ids is None else np.ascontiguousarray(hull_ids, dtype=np.int32)
    J = 0
    jp = None
    if joints is not None:
        joints = np.ascontiguousarray(joints, dtype=xo.JOINT_DTYPE).reshape(n_worlds, -1)
        J = joints.shape[1]
        jp = C.c_void_p(joints.ctypes.data)
    ctr = np.zeros(3, dtype=np.uint64)
    f32 = rigids38.dtype == np.float32
    fn = _L.xo_step_worlds_f32 if f32 else _L.xo_step_worlds_f64
    fn(C.c_void_p(rigids38.ctypes.data), n_worlds, B, C.cast(harr, _VP), len(hulls),
       None if ids is None else C.c_void_p(ids.ctypes.data), jp, J, compliance, friction, 1 if body_body else 0,
       np.float32(dt) if f32 else float(dt), substeps, frames, threads, C.c_void_p(ctr.ctypes.data))
    return {"constraints": int(ctr[0]), "pairs_tested": int(ctr[1]), "manifolds": int(ctr[2])}


PHASES = ("integrate", "ground", "broadphase", "face_queries", "edge_query", "manifold", "sweep", "derive", "velocity")
_L.xo_count_flops_worlds_f32.restype = C.c_uint64
_L.xo_count_flops_worlds_f32.argtypes = [_VP, C.c_int64, C.c_int, _VP, C.c_int, _VP, _VP, C.c_int, C.c_double, C.c_double,
                                         C.c_int, C.c_float, C.c_int, C.c_int, _VP]


def count_flops_worlds(rigids38_f32, B, hulls, dt, substeps, frames, hull_ids=None, joints=None, compliance=1e-6, friction=1.0,
                       body_body=True):
    """Algorithmic flops of the specification (oracle instantiated with the counting scalar, oracle/xo_count.hpp): steps a
    COPY of the given fp32 worlds for `frames` frames and returns {"constraints", "body_substeps", "flops": {phase: {add,
    mul, div, sqrt}}, "total"} with add = mul = div = sqrt = 1 flop."""
    r = np.ascontiguousarray(rigids38_f32, dtype=np.float32)
    n_worlds = r.size // (38 * B)
    harr, _keep = _hull_array(hulls)
    ids = None if hull_ids is None else np.ascontiguousarray(hull_ids, dtype=np.int32)
    J, jp = 0, None
    if joints is not None:
        joints = np.ascontiguousarray(joints, dtype=xo.JOINT_DTYPE).reshape(n_worlds, -1)
        J, jp = joints.shape[1], C.c_void_p(joints.ctypes.data)
    out = np.zeros((len(PHASES), 4), dtype=np.uint64)
    cons = _L.xo_count_flops_worlds_f32(C.c_void_p(r.ctypes.data), n_worlds, B, C.cast(harr, _VP), len(hulls),
                                        None if ids is None else C.c_void_p(ids.ctypes.data), jp, J, compliance, friction,
                                        1 if body_body else 0, np.float32(dt), substeps, frames, C.c_void_p(out.ctypes.data))
    flops = {p: dict(zip(("add", "mul", "div", "sqrt"), map(int, out[k]))) for k, p in enumerate(PHASES)}
    return {"constraints": int(cons), "body_substeps": n_worlds * B * substeps * frames, "flops": flops, "total": int(out.sum())}


def sat_manifold(fa7, fb7, hull_a, hull_b, dtype=np.float32):
    fa = np.ascontiguousarray(fa7, dtype=dtype); fb = np.ascontiguousarray(fb7, dtype=dtype)
    pts = np.zeros((4, 3), dtype=dtype); tg = np.zeros((4, 3), dtype=dtype)
    feat = C.c_int32(0); inc = C.c_int32(0)
    ka, kb = xo.Hull.of(hull_a), xo.Hull.of(hull_b)
    ha, hb = ka.as_c(), kb.as_c()
    fn = _L.xo_sat_manifold_f32 if dtype == np.float32 else _L.xo_sat_manifold_f64
    n = fn(C.c_void_p(fa.ctypes.data), C.c_void_p(fb.ctypes.data), C.byref(ha), C.byref(hb), C.c_void_p(pts.ctypes.data),
           C.c_void_p(tg.ctypes.data), C.byref(feat), C.byref(inc))
    return pts[:n], tg[:n], feat.value, inc.value
