"""Host-side mirror of the reference's `collision` module (src/collision.rs); each function runs the
corresponding CUDA kernel through the C ABI."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _abi
from .constraint import Constraint
from .frame import Frame
from .geometry import Polytope
from .rigid import Rigid


def ground(rigid: Rigid, past: Frame, polytope: Polytope) -> list[Constraint]:
    """collision::ground (src/collision.rs:13-35)"""
    nv = len(polytope.vertices)
    out = np.zeros(nv, dtype=_abi.CONSTRAINT_DTYPE)
    count = np.zeros(1, dtype=np.int32)
    past_rec = np.array([past.as_record()], dtype=_abi.FRAME_DTYPE)
    c = polytope.as_c()
    _abi.check(_abi.load().xpbd_ground_batch(C.c_void_p(rigid.rec.ctypes.data), _abi.ptr(past_rec), 1, C.byref(c),
                                             _abi.ptr(out), _abi.ptr(count)))
    return [Constraint.from_record(out[i]) for i in range(int(count[0]))]


def _frames(frames):
    a = np.array([frames[0].as_record()], dtype=_abi.FRAME_DTYPE)
    b = np.array([frames[1].as_record()], dtype=_abi.FRAME_DTYPE)
    return a, b


def face_axes_separation(frames, polytopes) -> tuple[float, int]:
    """collision::face_axes_separation (src/collision.rs:123-149)"""
    fa, fb = _frames(frames)
    dist = np.zeros(1); face = np.zeros(1, dtype=np.int32)
    ca, cb = polytopes[0].as_c(), polytopes[1].as_c()
    _abi.check(_abi.load().xpbd_face_axes_separation_batch(_abi.ptr(fa), _abi.ptr(fb), 1, C.byref(ca), C.byref(cb),
                                                           _abi.ptr(dist), _abi.ptr(face)))
    return float(dist[0]), int(face[0])


def edge_axes_separation(frames, polytopes) -> tuple[float, tuple[int, int]]:
    """collision::edge_axes_separation (src/collision.rs:151-197)"""
    fa, fb = _frames(frames)
    dist = np.zeros(1); edges = np.zeros((1, 2), dtype=np.int32)
    ca, cb = polytopes[0].as_c(), polytopes[1].as_c()
    _abi.check(_abi.load().xpbd_edge_axes_separation_batch(_abi.ptr(fa), _abi.ptr(fb), 1, C.byref(ca), C.byref(cb),
                                                           _abi.ptr(dist), _abi.ptr(edges)))
    return float(dist[0]), (int(edges[0, 0]), int(edges[0, 1]))


def sat(frames, polytopes) -> tuple[bool, list[Constraint]]:
    """collision::sat (src/collision.rs:37-121) finished per its commented plan: returns
    (separated, contact constraints). `separated` keeps the reference's sense: True = no contact."""
    fa, fb = _frames(frames)
    out = np.zeros(_abi.XPBD_MAX_MANIFOLD, dtype=_abi.CONSTRAINT_DTYPE)
    count = np.zeros(1, dtype=np.int32); feature = np.zeros(1, dtype=np.int32)
    ca, cb = polytopes[0].as_c(), polytopes[1].as_c()
    _abi.check(_abi.load().xpbd_sat_manifold_batch(_abi.ptr(fa), _abi.ptr(fb), 1, C.byref(ca), C.byref(cb),
                                                   _abi.ptr(out), _abi.ptr(count), _abi.ptr(feature)))
    return int(feature[0]) == 0, [Constraint.from_record(out[i]) for i in range(int(count[0]))]


SAT_SCHEDULES = {"warp": 0, "dual": 1, "thread": 2}


def sat_feature_batch(frames_a: np.ndarray, frames_b: np.ndarray, hull_a: Polytope, hull_b: Polytope, schedule: str = "warp",
                      cull: bool = True) -> tuple[np.ndarray, np.ndarray]:
    """The step's SAT stage (face query A, face query B, edge query, feature choice) on n synthetic candidate pairs,
    through the step's own kernels (xpbd_sat_feature_batch).  frames_*: FRAME_DTYPE arrays.  Returns (feature, index)."""
    fa = np.ascontiguousarray(frames_a, dtype=_abi.FRAME_DTYPE)
    fb = np.ascontiguousarray(frames_b, dtype=_abi.FRAME_DTYPE)
    assert fa.shape == fb.shape and fa.ndim == 1
    feature = np.zeros(len(fa), dtype=np.int32); index = np.zeros(len(fa), dtype=np.int32)
    ca, cb = hull_a.as_c(), hull_b.as_c()
    _abi.check(_abi.load().xpbd_sat_feature_batch(_abi.ptr(fa), _abi.ptr(fb), len(fa), C.byref(ca), C.byref(cb),
                                                  SAT_SCHEDULES[schedule], 1 if cull else 0, _abi.ptr(feature), _abi.ptr(index)))
    return feature, index
