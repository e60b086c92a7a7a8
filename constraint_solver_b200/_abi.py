"""ctypes binding of libxpbd_b200.so (include/xpbd_b200.h).

This is the only place Python touches the product library.  There is no CPU fallback: if the
shared library is missing, loading fails loudly; if no CUDA device is present, every compute
entry point returns XPBD_ERR_CUDA and `check` raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# XPBD_B200_LIB: another build of the same library (tuning variants under lib/variants/, the -DXPBD_CULL_STATS debug build)
LIB_PATH = os.environ.get("XPBD_B200_LIB") or os.path.join(_HERE, "lib", "libxpbd_b200.so")

XPBD_OK = 0
XPBD_ERR_INVALID = -1
XPBD_ERR_CUDA = -2
XPBD_ERR_CAPACITY = -3
XPBD_ERR_SINGULAR = -4
XPBD_ERR_STATE = -5
XPBD_MAX_MANIFOLD = 4
XPBD_STATIC_MAX_BODIES = 4096        # limits of xpbd_solve_batch_static (include/xpbd_b200.h)
XPBD_STATIC_MAX_CONSTRAINTS = 32768

# numpy mirrors of the C records (all C-contiguous, no padding surprises: checked below)
RIGID_DTYPE = np.dtype(
    [
        ("inverse_mass", "f8"),
        ("inverse_inertia", "f8", (9,)),
        ("external_force", "f8", (3,)),
        ("internal_force", "f8", (3,)),
        ("external_torque", "f8", (3,)),
        ("internal_torque", "f8", (3,)),
        ("velocity", "f8", (3,)),
        ("angular_velocity", "f8", (3,)),
        ("center_of_mass", "f8", (3,)),
        ("position", "f8", (3,)),
        ("rotation", "f8", (4,)),  # v.x v.y v.z s
    ]
)
assert RIGID_DTYPE.itemsize == 38 * 8
FRAME_DTYPE = np.dtype([("position", "f8", (3,)), ("rotation", "f8", (4,))])
assert FRAME_DTYPE.itemsize == 56
CONSTRAINT_DTYPE = np.dtype(
    [("rigid", "i4"), ("rigid_b", "i4"), ("contact0", "f8", (3,)), ("contact1", "f8", (3,)), ("distance", "f8")]
)
assert CONSTRAINT_DTYPE.itemsize == 64
JOINT_DTYPE = np.dtype(
    [("body_a", "i4"), ("body_b", "i4"), ("anchor_a", "f8", (3,)), ("anchor_b", "f8", (3,)), ("distance", "f8")]
)
assert JOINT_DTYPE.itemsize == 64
METRICS_DTYPE = np.dtype([("mass", "f8"), ("volume", "f8"), ("center_of_mass", "f8", (3,)), ("inertia_tensor", "f8", (9,))])
assert METRICS_DTYPE.itemsize == 14 * 8


class PolytopeC(C.Structure):
    _fields_ = [
        ("vertices", C.POINTER(C.c_double)),
        ("n_vertices", C.c_int32),
        ("edges", C.POINTER(C.c_int32)),
        ("n_edges", C.c_int32),
        ("face_offsets", C.POINTER(C.c_int32)),
        ("face_indices", C.POINTER(C.c_int32)),
        ("n_faces", C.c_int32),
        ("centroid", C.c_double * 3),
    ]


class BatchDescC(C.Structure):
    _fields_ = [
        ("device", C.c_int32),
        ("n_worlds", C.c_int64),
        ("bodies_per_world", C.c_int32),
        ("joints_per_world", C.c_int32),
        ("body_body_contacts", C.c_int32),
        ("compliance", C.c_double),
        ("friction", C.c_double),
    ]


class StepStatsC(C.Structure):
    _fields_ = [
        ("body_substeps", C.c_uint64),
        ("constraints", C.c_uint64),
        ("overflow_worlds", C.c_uint64),
        ("elapsed_ms", C.c_float),
        ("kernel_launches", C.c_int32),
        ("stage_ms", C.c_float * 4),
        ("candidate_pairs", C.c_uint64),
        ("sub_batches", C.c_int32),
        ("nan_worlds", C.c_int32),
    ]


class XpbdError(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__(f"xpbd_b200 error {status}: {message}")
        self.status = status


_VP = C.c_void_p
_PROTOTYPES = {
    # name: (restype, argtypes)
    "xpbd_last_error": (C.c_char_p, []),
    "xpbd_version": (C.c_char_p, []),
    "xpbd_device_count": (C.c_int, [C.POINTER(C.c_int32)]),
    "xpbd_polytope_rigid_metrics": (C.c_int, [C.POINTER(PolytopeC), C.c_double, _VP]),
    "xpbd_rigid_new": (C.c_int, [_VP, _VP]),
    "xpbd_polytope_planes": (C.c_int, [C.POINTER(PolytopeC), _VP]),
    "xpbd_quickhull": (C.c_int, [_VP, C.c_int32, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP, _VP]),
    "xpbd_batch_create": (C.c_int, [C.POINTER(BatchDescC), C.POINTER(_VP)]),
    "xpbd_batch_destroy": (C.c_int, [_VP]),
    "xpbd_batch_add_hull": (C.c_int, [_VP, C.POINTER(PolytopeC), C.POINTER(C.c_int32)]),
    "xpbd_batch_set_hulls": (C.c_int, [_VP, C.c_int64, C.c_int64, _VP]),
    "xpbd_batch_set_joints": (C.c_int, [_VP, C.c_int64, C.c_int64, _VP]),
    "xpbd_batch_set_stream": (C.c_int, [_VP, _VP]),
    "xpbd_batch_set_sub_batches": (C.c_int, [_VP, C.c_int32]),
    "xpbd_batch_upload": (C.c_int, [_VP, C.c_int64, C.c_int64, _VP]),
    "xpbd_batch_download": (C.c_int, [_VP, C.c_int64, C.c_int64, _VP]),
    "xpbd_batch_upload_broadcast": (C.c_int, [_VP, _VP]),
    "xpbd_batch_download_dynamic_f32": (C.c_int, [_VP, C.c_int64, C.c_int64, _VP]),
    "xpbd_batch_upload_dynamic_f32": (C.c_int, [_VP, C.c_int64, C.c_int64, _VP]),
    "xpbd_batch_step": (C.c_int, [_VP, C.c_double, C.c_int32, C.c_int32]),
    "xpbd_batch_step_host_f32": (C.c_int, [_VP, C.c_double, C.c_int32, C.c_int32, _VP, _VP]),
    "xpbd_batch_step_host_rigid": (C.c_int, [_VP, C.c_double, C.c_int32, C.c_int32, _VP, _VP]),
    "xpbd_batch_sync": (C.c_int, [_VP]),
    "xpbd_batch_last_stats": (C.c_int, [_VP, C.POINTER(StepStatsC)]),
    "xpbd_solver_step": (C.c_int, [_VP, C.POINTER(PolytopeC), C.c_double, C.c_int32]),
    "xpbd_world_integrate": (C.c_int, [_VP, _VP, C.c_double, C.POINTER(PolytopeC), C.POINTER(PolytopeC)]),
    "xpbd_rigid_integrate_batch": (C.c_int, [_VP, C.c_int64, C.c_double]),
    "xpbd_rigid_derive_batch": (C.c_int, [_VP, _VP, C.c_int64, C.c_double]),
    "xpbd_rigid_apply_impulse_batch": (C.c_int, [_VP, _VP, C.c_int64]),
    "xpbd_rigid_frame_batch": (C.c_int, [_VP, C.c_int64, _VP]),
    "xpbd_ground_batch": (C.c_int, [_VP, _VP, C.c_int64, C.POINTER(PolytopeC), _VP, _VP]),
    "xpbd_solve_batch": (C.c_int, [_VP, C.c_int64, _VP, _VP, C.c_int32, C.c_double, C.c_double]),
    "xpbd_solve_batch_static": (C.c_int, [_VP, C.c_int64, _VP, _VP, C.c_int32, C.c_double, C.c_double]),
    "xpbd_face_axes_separation_batch": (C.c_int, [_VP, _VP, C.c_int64, C.POINTER(PolytopeC), C.POINTER(PolytopeC), _VP, _VP]),
    "xpbd_edge_axes_separation_batch": (C.c_int, [_VP, _VP, C.c_int64, C.POINTER(PolytopeC), C.POINTER(PolytopeC), _VP, _VP]),
    "xpbd_sat_manifold_batch": (C.c_int, [_VP, _VP, C.c_int64, C.POINTER(PolytopeC), C.POINTER(PolytopeC), _VP, _VP, _VP]),
    "xpbd_sat_feature_batch": (C.c_int, [_VP, _VP, C.c_int64, C.POINTER(PolytopeC), C.POINTER(PolytopeC), C.c_int32, C.c_int32, _VP, _VP]),
    "xpbd_fp32_peak_microbench": (C.c_int, [C.c_int32, C.c_int32, C.POINTER(C.c_double)]),
}
EXPORTED_SYMBOLS = tuple(_PROTOTYPES)

_lib = None


def load() -> C.CDLL:
    """Load libxpbd_b200.so; raises if it has not been built (run `make lib` or __graft_entry__.build())."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise FileNotFoundError(
                f"{LIB_PATH} is missing: build it with `make lib` (nvcc, sm_100a). "
                "constraint_solver_b200 has no CPU or PyTorch fallback."
            )
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in _PROTOTYPES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def check(status: int) -> None:
    if status != XPBD_OK:
        raise XpbdError(status, load().xpbd_last_error().decode("utf-8", "replace"))


def ptr(a: np.ndarray) -> C.c_void_p:
    assert a.flags["C_CONTIGUOUS"], "array must be C-contiguous"
    return C.c_void_p(a.ctypes.data)
