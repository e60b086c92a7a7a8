"""Host-side mirror of `Rigid` (src/rigid.rs:7-124).

The record is the reference's own AoS f64 layout (include/xpbd_b200.h xpbd_rigid); every method
that does arithmetic runs the corresponding CUDA kernel through the C ABI (fp32 on the device).
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _abi
from .frame import Frame
from .geometry import RigidMetrics

_FIELDS = ("inverse_mass", "inverse_inertia", "external_force", "internal_force", "external_torque",
           "internal_torque", "velocity", "angular_velocity", "center_of_mass", "position", "rotation")


class Rigid:
    __slots__ = ("rec", "color")

    def __init__(self, metrics: RigidMetrics | None = None, rec: np.ndarray | None = None):
        """Rigid::new (src/rigid.rs:53-71). Raises XpbdError(XPBD_ERR_SINGULAR) where the reference panics."""
        self.color = None  # src/rigid.rs:49; host-only
        if rec is not None:
            self.rec = np.array(rec, dtype=_abi.RIGID_DTYPE).reshape(())
            return
        self.rec = np.zeros((), dtype=_abi.RIGID_DTYPE)
        m = metrics.as_record()
        _abi.check(_abi.load().xpbd_rigid_new(C.c_void_p(m.ctypes.data), C.c_void_p(self.rec.ctypes.data)))

    def __getattr__(self, name):
        if name in _FIELDS:
            v = self.rec[name]
            return float(v) if v.ndim == 0 else v
        raise AttributeError(name)

    def __setattr__(self, name, value):
        if name in _FIELDS:
            self.rec[name] = value
        else:
            object.__setattr__(self, name, value)

    def copy(self) -> "Rigid":  # Rigid is Copy in the reference
        r = Rigid(rec=self.rec.copy())
        r.color = self.color
        return r

    def frame(self) -> Frame:  # src/rigid.rs:75-80
        out = np.zeros(1, dtype=_abi.FRAME_DTYPE)
        _abi.check(_abi.load().xpbd_rigid_frame_batch(C.c_void_p(self.rec.ctypes.data), 1, _abi.ptr(out)))
        return Frame.from_record(out[0])

    def integrate(self, dt: float) -> None:  # src/rigid.rs:82-99
        _abi.check(_abi.load().xpbd_rigid_integrate_batch(C.c_void_p(self.rec.ctypes.data), 1, float(dt)))

    def derive(self, position, rotation, dt: float) -> None:  # src/rigid.rs:101-109
        past = np.concatenate([np.asarray(position, dtype=np.float64), np.asarray(rotation, dtype=np.float64)])
        _abi.check(_abi.load().xpbd_rigid_derive_batch(C.c_void_p(self.rec.ctypes.data), _abi.ptr(past), 1, float(dt)))

    def apply_impulse(self, impulse, point) -> None:  # src/rigid.rs:113-123
        ip = np.concatenate([np.asarray(impulse, dtype=np.float64), np.asarray(point, dtype=np.float64)])
        _abi.check(_abi.load().xpbd_rigid_apply_impulse_batch(C.c_void_p(self.rec.ctypes.data), _abi.ptr(ip), 1))
