"""Host-side mirror of the reference's `geometry` module (src/geometry.rs, src/geometry/integrate.rs).

Polytope holds the hull arrays that get registered in the device hull table; the setup-time
math (face planes, Mirtich mass properties) runs in the native library in f64.
"""
from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass

import numpy as np

from . import _abi


@dataclass
class Plane:
    """src/geometry.rs:8-12"""

    normal: np.ndarray
    displacement: float

    def distance(self, point) -> float:  # src/geometry.rs:39-41
        return float(np.dot(self.normal, point) - self.displacement)

    def constant(self) -> float:  # src/geometry.rs:50-52
        return -self.displacement

    def flip(self) -> "Plane":  # src/geometry.rs:55-60
        return Plane(-self.normal, -self.displacement)


@dataclass
class RigidMetrics:
    """src/geometry/integrate.rs:18-24"""

    mass: float
    volume: float
    center_of_mass: np.ndarray
    inertia_tensor: np.ndarray  # 3x3, inertia_tensor[col, row] as cgmath stores it

    def as_record(self) -> np.ndarray:
        r = np.zeros((), dtype=_abi.METRICS_DTYPE)
        r["mass"], r["volume"] = self.mass, self.volume
        r["center_of_mass"] = self.center_of_mass
        r["inertia_tensor"] = np.asarray(self.inertia_tensor, dtype=np.float64).reshape(9)
        return r


class Polytope:
    """src/geometry.rs:81-93: vertices, edges (vertex index pairs), faces (convex polygons), centroid."""

    def __init__(self, vertices, edges, faces, centroid):
        self.vertices = np.ascontiguousarray(vertices, dtype=np.float64).reshape(-1, 3)
        self.edges = np.ascontiguousarray(edges, dtype=np.int32).reshape(-1, 2)
        self.faces = [list(map(int, f)) for f in faces]
        self.centroid = np.ascontiguousarray(centroid, dtype=np.float64).reshape(3)
        self._face_offsets = np.zeros(len(self.faces) + 1, dtype=np.int32)
        self._face_offsets[1:] = np.cumsum([len(f) for f in self.faces])
        self._face_indices = np.ascontiguousarray([v for f in self.faces for v in f], dtype=np.int32)

    # -- the three built-in hulls ---------------------------------------------------------
    @staticmethod
    def new_tetrahedron() -> "Polytope":  # src/geometry.rs:97-109
        return Polytope(
            [(0, 0, 0), (1, 0, 0), (0, 1, 0), (0, 0, 1)],
            [(0, 1), (0, 2), (0, 3), (1, 2), (1, 3), (2, 3)],
            [[0, 3, 2], [3, 0, 1], [2, 1, 0], [1, 2, 3]],
            (0.25, 0.25, 0.25),
        )

    @staticmethod
    def new_cube() -> "Polytope":  # src/geometry.rs:113-149
        return Polytope(
            [(0, 0, 0), (1, 0, 0), (0, 1, 0), (1, 1, 0), (0, 0, 1), (1, 0, 1), (0, 1, 1), (1, 1, 1)],
            [(0, 1), (1, 3), (3, 2), (2, 0), (4, 5), (5, 7), (7, 6), (6, 4), (0, 4), (1, 5), (3, 7), (2, 6)],
            [[0, 2, 3, 1], [4, 5, 7, 6], [4, 0, 1, 5], [5, 1, 3, 7], [7, 3, 2, 6], [6, 2, 0, 4]],
            (0.5, 0.5, 0.5),
        )

    @staticmethod
    def new_icosahedron() -> "Polytope":  # src/geometry.rs:153-230
        phi = (1.0 + math.sqrt(5.0)) / 2.0
        mag = math.sqrt(phi * phi + 1.0)
        a, b = phi / mag, 1.0 / mag
        verts = [(a, b, 0), (a, -b, 0), (-a, b, 0), (-a, -b, 0), (0, a, b), (0, a, -b),
                 (0, -a, b), (0, -a, -b), (b, 0, a), (-b, 0, a), (b, 0, -a), (-b, 0, -a)]
        edges = [(8, 9), (8, 0), (8, 1), (1, 0), (9, 2), (9, 3), (2, 3), (2, 5), (2, 11), (5, 11),
                 (3, 7), (3, 11), (7, 11), (0, 5), (1, 7), (4, 5), (4, 0), (4, 8), (4, 9), (4, 2),
                 (6, 9), (6, 8), (6, 1), (6, 7), (6, 3), (10, 1), (10, 0), (10, 5), (10, 11), (10, 7)]
        faces = [[0, 5, 4], [2, 4, 5], [1, 6, 7], [3, 7, 6], [1, 0, 8], [0, 1, 10], [2, 3, 9],
                 [3, 2, 11], [4, 9, 8], [6, 8, 9], [5, 10, 11], [7, 11, 10], [0, 4, 8], [0, 10, 5],
                 [2, 9, 4], [2, 5, 11], [1, 8, 6], [1, 7, 10], [3, 6, 9], [3, 11, 7]]
        return Polytope(verts, edges, faces, (0.0, 0.0, 0.0))

    @staticmethod
    def from_points(points) -> "Polytope":
        """QuickHull of a point cloud (README.md:20), computed by the native library."""
        pts = np.ascontiguousarray(points, dtype=np.float64).reshape(-1, 3)
        n = len(pts)
        cap_f = max(4, 2 * n)
        verts = np.zeros((n, 3)); edges = np.zeros((3 * n + 6, 2), dtype=np.int32)
        foff = np.zeros(cap_f + 1, dtype=np.int32); fidx = np.zeros(6 * n + 12, dtype=np.int32)
        nv = C.c_int32(n); ne = C.c_int32(len(edges)); nf = C.c_int32(cap_f); nfi = C.c_int32(len(fidx))
        centroid = np.zeros(3)
        _abi.check(_abi.load().xpbd_quickhull(_abi.ptr(pts), n, _abi.ptr(verts), C.addressof(nv), _abi.ptr(edges),
                                               C.addressof(ne), _abi.ptr(foff), _abi.ptr(fidx), C.addressof(nf),
                                               C.addressof(nfi), _abi.ptr(centroid)))
        faces = [fidx[foff[f]:foff[f + 1]].tolist() for f in range(nf.value)]
        return Polytope(verts[: nv.value], edges[: ne.value], faces, centroid)

    # -- ABI view ----------------------------------------------------------------------------
    def as_c(self) -> _abi.PolytopeC:
        p = _abi.PolytopeC()
        p.vertices = self.vertices.ctypes.data_as(C.POINTER(C.c_double))
        p.n_vertices = len(self.vertices)
        p.edges = self.edges.ctypes.data_as(C.POINTER(C.c_int32))
        p.n_edges = len(self.edges)
        p.face_offsets = self._face_offsets.ctypes.data_as(C.POINTER(C.c_int32))
        p.face_indices = self._face_indices.ctypes.data_as(C.POINTER(C.c_int32))
        p.n_faces = len(self.faces)
        p.centroid[:] = self.centroid.tolist()
        return p

    # -- reference methods -------------------------------------------------------------------
    def planes(self) -> list[Plane]:  # src/geometry.rs:256-258
        out = np.zeros((len(self.faces), 4))
        c = self.as_c()
        _abi.check(_abi.load().xpbd_polytope_planes(C.byref(c), _abi.ptr(out)))
        return [Plane(out[i, :3].copy(), float(out[i, 3])) for i in range(len(self.faces))]

    def plane(self, i: int) -> Plane:  # src/geometry.rs:262-271
        return self.planes()[i]

    def rigid_metrics(self, density: float) -> RigidMetrics:  # src/geometry.rs:291
        rec = np.zeros((), dtype=_abi.METRICS_DTYPE)
        c = self.as_c()
        _abi.check(_abi.load().xpbd_polytope_rigid_metrics(C.byref(c), float(density), C.c_void_p(rec.ctypes.data)))
        return RigidMetrics(float(rec["mass"]), float(rec["volume"]), rec["center_of_mass"].copy(),
                            rec["inertia_tensor"].reshape(3, 3).copy())

    def scaled(self, k: float) -> "Polytope":  # `f64 * Polytope`, src/geometry.rs:296-307
        return Polytope(k * self.vertices, self.edges, self.faces, k * self.centroid)

    def __rmul__(self, k: float) -> "Polytope":
        return self.scaled(float(k))
