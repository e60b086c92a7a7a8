"""Host-side mirror of `World` (src/world.rs:5-43) and its batched form `Worlds`.

`World` is the reference's two-body scene, stepped through the single-call drop-in.
`Worlds` is the B200-native batch: n_worlds independent worlds resident in HBM as fp32 SoA,
sharded across GPUs by world index (one `Worlds` per rank/GPU), stepped by one fused launch.
"""
from __future__ import annotations

import ctypes as C
import math

import numpy as np

from . import _abi
from .geometry import Polytope
from .rigid import Rigid

SUBSTEPS = 25  # src/world.rs:41-42


def _quat_from_euler_deg(x: float, y: float, z: float) -> np.ndarray:
    """cgmath 0.18 Quaternion::from(Euler{x,y,z}) (src/world.rs:28). Returns v.x v.y v.z s."""
    rx, ry, rz = (a * (math.pi / 180.0) for a in (x, y, z))
    sx, cx = math.sin(rx * 0.5), math.cos(rx * 0.5)
    sy, cy = math.sin(ry * 0.5), math.cos(ry * 0.5)
    sz, cz = math.sin(rz * 0.5), math.cos(rz * 0.5)
    s = -sx * sy * sz + cx * cy * cz
    qx = sx * cy * cz + sy * sz * cx
    qy = -sx * sz * cy + sy * cx * cz
    qz = sx * sy * cz + sz * cx * cy
    return np.array([qx, qy, qz, s])


def shard_range(n_worlds_total: int, rank: int, world_size: int) -> tuple[int, int]:
    """Contiguous world-index range [first, first + count) owned by `rank`.

    Worlds never interact (src/world.rs:41-42 steps even the two bodies of one world independently), so the
    batch shards by world index with no collective on the step path.  Remainders go to the lowest ranks."""
    if world_size <= 0 or not (0 <= rank < world_size) or n_worlds_total < 0:
        raise ValueError("bad shard arguments")
    base, rem = divmod(n_worlds_total, world_size)
    first = rank * base + min(rank, rem)
    return first, base + (1 if rank < rem else 0)


class World:
    """src/world.rs:5-9"""

    def __init__(self, p1: Polytope, p2: Polytope):  # World::new, src/world.rs:12-31
        a = Rigid(p1.rigid_metrics(0.1))
        b = Rigid(p2.rigid_metrics(5.0))
        a.position = (0.0, 0.0, 4.0)
        a.velocity = (0.0, 2.5, 0.0)
        a.angular_velocity = (-4.0, 1.0, 0.0)
        a.external_force = (0.0, 0.0, -2.0)
        b.position = (4.0, 0.0, 4.0)
        b.velocity = (0.0, 0.0, 7.0)
        b.angular_velocity = (-5.0, 5.0, 0.0)
        b.external_force = (0.0, 0.0, -2.0)
        b.rotation = _quat_from_euler_deg(10.0, 15.0, 5.0)
        self.a, self.b = a, b

    def copy(self) -> "World":  # World is Copy (src/world.rs:5); the app's rewind buffer relies on it
        w = World.__new__(World)
        w.a, w.b = self.a.copy(), self.b.copy()
        return w

    def integrate(self, dt: float, p1: Polytope, p2: Polytope, debug=None) -> None:
        """World::integrate (src/world.rs:34-43): both bodies against p1 (reference quirk); debug unused."""
        c1, c2 = p1.as_c(), p2.as_c()
        _abi.check(_abi.load().xpbd_world_integrate(C.c_void_p(self.a.rec.ctypes.data), C.c_void_p(self.b.rec.ctypes.data),
                                                    float(dt), C.byref(c1), C.byref(c2)))


class Worlds:
    """A batch of independent worlds on one GPU (opaque xpbd_batch handle)."""

    def __init__(self, n_worlds: int, bodies_per_world: int = 1, *, device: int = 0, joints_per_world: int = 0,
                 body_body_contacts: bool = False, compliance: float = 1e-6, friction: float = 1.0):
        self._lib = _abi.load()
        self.n_worlds, self.bodies_per_world, self.joints_per_world = int(n_worlds), int(bodies_per_world), int(joints_per_world)
        desc = _abi.BatchDescC(int(device), int(n_worlds), int(bodies_per_world), int(joints_per_world),
                               1 if body_body_contacts else 0, float(compliance), float(friction))
        self._h = C.c_void_p()
        _abi.check(self._lib.xpbd_batch_create(C.byref(desc), C.byref(self._h)))
        self._keep = []  # keeps hull arrays alive while the C view is in use

    def close(self) -> None:
        if getattr(self, "_h", None) and self._h.value:
            self._lib.xpbd_batch_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    @property
    def n_bodies(self) -> int:
        return self.n_worlds * self.bodies_per_world

    def add_hull(self, polytope: Polytope) -> int:
        hid = C.c_int32(-1)
        c = polytope.as_c()
        _abi.check(self._lib.xpbd_batch_add_hull(self._h, C.byref(c), C.byref(hid)))
        return hid.value

    def set_hulls(self, ids, first_world: int = 0) -> None:
        ids = np.ascontiguousarray(ids, dtype=np.int32).reshape(-1, self.bodies_per_world)
        _abi.check(self._lib.xpbd_batch_set_hulls(self._h, first_world, len(ids), _abi.ptr(ids)))

    def set_joints(self, joints: np.ndarray, first_world: int = 0) -> None:
        joints = np.ascontiguousarray(joints, dtype=_abi.JOINT_DTYPE).reshape(-1, self.joints_per_world)
        _abi.check(self._lib.xpbd_batch_set_joints(self._h, first_world, len(joints), _abi.ptr(joints)))

    def set_stream(self, cuda_stream: int) -> None:
        _abi.check(self._lib.xpbd_batch_set_stream(self._h, C.c_void_p(cuda_stream)))

    def set_sub_batches(self, count: int) -> None:
        """World ranges stepped concurrently by the staged multi-body pipeline (0 = automatic, 1 = a single chain of
        launches; results do not depend on it)."""
        _abi.check(self._lib.xpbd_batch_set_sub_batches(self._h, int(count)))

    def upload(self, rigids: np.ndarray, first_world: int = 0) -> None:
        rigids = np.ascontiguousarray(rigids, dtype=_abi.RIGID_DTYPE).reshape(-1, self.bodies_per_world)
        _abi.check(self._lib.xpbd_batch_upload(self._h, first_world, len(rigids), _abi.ptr(rigids)))

    def upload_broadcast(self, world_bodies: np.ndarray) -> None:
        wb = np.ascontiguousarray(world_bodies, dtype=_abi.RIGID_DTYPE).reshape(self.bodies_per_world)
        _abi.check(self._lib.xpbd_batch_upload_broadcast(self._h, _abi.ptr(wb)))

    def download(self, first_world: int = 0, n_worlds: int | None = None) -> np.ndarray:
        n = self.n_worlds - first_world if n_worlds is None else n_worlds
        out = np.zeros((n, self.bodies_per_world), dtype=_abi.RIGID_DTYPE)
        _abi.check(self._lib.xpbd_batch_download(self._h, first_world, n, _abi.ptr(out)))
        return out

    def download_dynamic(self, first_world: int = 0, n_worlds: int | None = None, out: np.ndarray | None = None) -> np.ndarray:
        n = self.n_worlds - first_world if n_worlds is None else n_worlds
        if out is None:
            out = np.zeros((n, self.bodies_per_world, 13), dtype=np.float32)
        _abi.check(self._lib.xpbd_batch_download_dynamic_f32(self._h, first_world, n, C.c_void_p(out.ctypes.data)))
        return out

    def upload_dynamic(self, dyn: np.ndarray, first_world: int = 0) -> None:
        dyn = np.ascontiguousarray(dyn, dtype=np.float32).reshape(-1, self.bodies_per_world, 13)
        _abi.check(self._lib.xpbd_batch_upload_dynamic_f32(self._h, first_world, len(dyn), _abi.ptr(dyn)))

    def upload_dynamic_ptr(self, host_ptr: int, first_world: int, n_worlds: int) -> None:
        _abi.check(self._lib.xpbd_batch_upload_dynamic_f32(self._h, first_world, n_worlds, C.c_void_p(host_ptr)))

    def download_dynamic_ptr(self, host_ptr: int, first_world: int, n_worlds: int) -> None:
        _abi.check(self._lib.xpbd_batch_download_dynamic_f32(self._h, first_world, n_worlds, C.c_void_p(host_ptr)))

    def integrate(self, dt: float, substep_count: int = SUBSTEPS, n_frames: int = 1, sync: bool = True) -> None:
        """World::integrate for every world: `n_frames` frames of `substep_count` substeps, one fused launch."""
        _abi.check(self._lib.xpbd_batch_step(self._h, float(dt), int(substep_count), int(n_frames)))
        if sync:
            self.sync()

    def integrate_host_ptr(self, dt: float, substep_count: int, n_frames: int, in_ptr: int, out_ptr: int, sync: bool = True) -> None:
        """World::integrate for callers whose state lives in host memory: packed dynamic state in (13 floats per body,
        0 = keep the device state), the same out (0 = none); one call, copies pipelined under the kernels per world range."""
        _abi.check(self._lib.xpbd_batch_step_host_f32(self._h, float(dt), int(substep_count), int(n_frames),
                                                      C.c_void_p(in_ptr or None), C.c_void_p(out_ptr or None)))
        if sync:
            self.sync()

    def integrate_host(self, dyn_in: np.ndarray | None, dt: float, substep_count: int = SUBSTEPS, n_frames: int = 1) -> np.ndarray:
        """Numpy form of integrate_host_ptr: returns the new packed dynamic state (n_worlds x bodies_per_world x 13 float32)."""
        out = np.empty((self.n_worlds, self.bodies_per_world, 13), dtype=np.float32)  # download_dynamic's shape
        if dyn_in is not None:
            dyn_in = np.ascontiguousarray(dyn_in, dtype=np.float32)
            assert dyn_in.size == self.n_bodies * 13
        self.integrate_host_ptr(dt, substep_count, n_frames, 0 if dyn_in is None else dyn_in.ctypes.data, out.ctypes.data)
        return out

    def integrate_host_rigid_ptr(self, dt: float, substep_count: int, n_frames: int, in_ptr: int, out_ptr: int, sync: bool = True) -> None:
        """World::integrate for callers holding the reference's own f64 records (xpbd_rigid, 304 B per body) in host memory:
        upload (0 = keep the device state), step, write the records back (0 = none) — xpbd_batch_step_host_rigid."""
        _abi.check(self._lib.xpbd_batch_step_host_rigid(self._h, float(dt), int(substep_count), int(n_frames),
                                                        C.c_void_p(in_ptr or None), C.c_void_p(out_ptr or None)))
        if sync:
            self.sync()

    def integrate_host_rigid(self, rigids: np.ndarray, dt: float, substep_count: int = SUBSTEPS, n_frames: int = 1) -> np.ndarray:
        """Numpy form: steps the given records (RIGID_DTYPE, n_worlds x bodies_per_world) and returns the new records."""
        rec = np.ascontiguousarray(rigids, dtype=_abi.RIGID_DTYPE)
        assert rec.size == self.n_bodies
        out = np.empty_like(rec)
        self.integrate_host_rigid_ptr(dt, substep_count, n_frames, rec.ctypes.data, out.ctypes.data)
        return out

    def sync(self) -> None:
        _abi.check(self._lib.xpbd_batch_sync(self._h))

    def stats(self) -> dict:
        s = _abi.StepStatsC()
        _abi.check(self._lib.xpbd_batch_last_stats(self._h, C.byref(s)))
        out = {f: getattr(s, f) for f, _ in s._fields_}
        out["stage_ms"] = [float(x) for x in s.stage_ms]
        return out
