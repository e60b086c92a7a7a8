"""Synthetic scenes of BASELINE.json's configs, as xpbd_rigid records.

C1  the reference's default cube: World::new body `a` (src/world.rs:13,16-20)
C2  2^20 single-cube worlds: identical copies of C1, or perturbed (Philox counter RNG, seed 0x5EED)
C3  10-cube stack per world, 20 substeps/frame
C4  64-body pile of mixed hulls with attachment/distance joints
"""
from __future__ import annotations

import numpy as np

from . import _abi
from .geometry import Polytope
from .rigid import Rigid
from .world import World

SEED = 0x5EED
FRAME_DT = 1.0 / 60.0  # src/app.rs:15


def default_world() -> World:
    """app.rs:29-30,48: World::new(cube, 0.5 * tetrahedron)"""
    return World(Polytope.new_cube(), 0.5 * Polytope.new_tetrahedron())


def cube_body_a() -> np.ndarray:
    return default_world().a.rec.copy()


def c2_identical(n_worlds: int) -> np.ndarray:
    out = np.empty((n_worlds, 1), dtype=_abi.RIGID_DTYPE)
    out[:] = cube_body_a()
    return out


def random_unit_quaternions(rng: np.random.Generator, n: int) -> np.ndarray:
    """Uniform on S^3 (Shoemake). Returned as v.x v.y v.z s."""
    u = rng.random((n, 3))
    a, b = np.sqrt(1.0 - u[:, 0]), np.sqrt(u[:, 0])
    t1, t2 = 2.0 * np.pi * u[:, 1], 2.0 * np.pi * u[:, 2]
    return np.stack([a * np.sin(t1), a * np.cos(t1), b * np.sin(t2), b * np.cos(t2)], axis=1)


def c2_perturbed(n_worlds: int, first_world: int = 0) -> np.ndarray:
    """z in U[2,6], v in U[-3,3]^3, w in U[-5,5]^3, uniform random orientation; keyed on world index so that
    a shard [first_world, first_world + n) gets the same worlds as the unsharded batch."""
    out = np.empty((n_worlds, 1), dtype=_abi.RIGID_DTYPE)
    out[:] = cube_body_a()
    # Philox is counter-based: advance by world index so shards are reproducible independently.
    bitgen = np.random.Philox(key=SEED)
    bitgen.advance(first_world * 4)  # one Philox counter step = 4 x 64 bits; 16 draws reserved per world (10 used)
    rng = np.random.Generator(bitgen)
    u = rng.random((n_worlds, 16))
    out["position"][:, 0, 2] = 2.0 + 4.0 * u[:, 0]
    out["velocity"][:, 0, :] = -3.0 + 6.0 * u[:, 1:4]
    out["angular_velocity"][:, 0, :] = -5.0 + 10.0 * u[:, 4:7]
    a, b = np.sqrt(1.0 - u[:, 7]), np.sqrt(u[:, 7])
    t1, t2 = 2.0 * np.pi * u[:, 8], 2.0 * np.pi * u[:, 9]
    out["rotation"][:, 0, :] = np.stack([a * np.sin(t1), a * np.cos(t1), b * np.sin(t2), b * np.cos(t2)], axis=1)
    return out


def cube_rigid(density: float = 0.1) -> Rigid:
    return Rigid(Polytope.new_cube().rigid_metrics(density))


def c3_stack(n_worlds: int, first_world: int = 0, n_cubes: int = 10, gap: float = 1e-3, jitter: float = 0.01) -> np.ndarray:
    """10 unit cubes (m = 0.1) per world stacked on the z axis, seeded +-jitter in x/y, `gap` between cubes,
    F_ext = (0, 0, -2) each."""
    base = cube_rigid(0.1)
    base.external_force = (0.0, 0.0, -2.0)
    out = np.empty((n_worlds, n_cubes), dtype=_abi.RIGID_DTYPE)
    out[:] = base.rec
    bitgen = np.random.Philox(key=SEED + 3)
    per_world = 4 * ((2 * n_cubes + 3) // 4)  # whole Philox counter steps per world
    bitgen.advance(first_world * per_world // 4)
    rng = np.random.Generator(bitgen)
    u = rng.random((n_worlds, per_world))[:, : 2 * n_cubes].reshape(n_worlds, n_cubes, 2)
    out["position"][:, :, 0:2] = -jitter + 2.0 * jitter * u
    out["position"][:, :, 2] = gap + np.arange(n_cubes)[None, :] * (1.0 + gap)
    return out


def c4_hulls() -> list[Polytope]:
    """Hull table of C4: the reference's three built-ins plus five QuickHull hulls of 8..32 seeded points on a
    sphere of radius 0.6 (QuickHull = xpbd_quickhull, row F3)."""
    hulls = [Polytope.new_cube(), Polytope.new_tetrahedron(), 0.6 * Polytope.new_icosahedron()]
    rng = np.random.Generator(np.random.Philox(key=SEED + 40))
    for n in (8, 12, 16, 24, 32):
        p = rng.normal(size=(n, 3))
        p *= 0.6 / np.linalg.norm(p, axis=1, keepdims=True)
        hulls.append(Polytope.from_points(p))
    return hulls


def c4_pile(n_worlds: int, first_world: int = 0, n_side: int = 4, joints_per_world: int = 32, spacing: float = 2.2,
            jitter: float = 0.1):
    """64-body pile per world: bodies drawn from c4_hulls() on an n_side^3 lattice (seeded jitter and orientation),
    falling onto the ground under a = -20 m/s^2, with `joints_per_world` distance joints (Constraint.distance,
    src/constraint.rs:9) between the centres of mass of x-neighbours at their initial separation.
    Returns (rigids [n_worlds, B], hulls, hull_ids [n_worlds, B] int32, joints [n_worlds, J])."""
    hulls = c4_hulls()
    B = n_side ** 3
    protos = []
    for k, h in enumerate(hulls):
        r = Rigid(h.rigid_metrics(0.1 if k == 0 else 0.3))
        r.external_force = (0.0, 0.0, -20.0 / r.rec["inverse_mass"])  # same acceleration as the reference's cube
        protos.append(r.rec.copy())
    bitgen = np.random.Philox(key=SEED + 4)
    per_world = 4 * ((B * 8 + 3) // 4)  # 8 draws per body, whole Philox counter steps
    bitgen.advance(first_world * per_world // 4)
    rng = np.random.Generator(bitgen)
    u = rng.random((n_worlds, per_world))[:, : B * 8].reshape(n_worlds, B, 8)
    hull_ids = np.minimum((u[:, :, 0] * len(hulls)).astype(np.int32), len(hulls) - 1)
    out = np.empty((n_worlds, B), dtype=_abi.RIGID_DTYPE)
    for k, p in enumerate(protos):
        out[hull_ids == k] = p
    ix, iy, iz = np.meshgrid(np.arange(n_side), np.arange(n_side), np.arange(n_side), indexing="ij")
    lattice = np.stack([ix.ravel(), iy.ravel(), iz.ravel()], axis=1).astype(np.float64)  # body b -> (x, y, z) cell, x-major
    centre = lattice * spacing + np.array([0.0, 0.0, 1.5])
    com_world = centre[None, :, :] + (-jitter + 2.0 * jitter * u[:, :, 1:4])
    out["position"] = com_world - out["center_of_mass"]
    a, b = np.sqrt(1.0 - u[:, :, 4]), np.sqrt(u[:, :, 4])
    t1, t2 = 2.0 * np.pi * u[:, :, 5], 2.0 * np.pi * u[:, :, 6]
    out["rotation"] = np.stack([a * np.sin(t1), a * np.cos(t1), b * np.sin(t2), b * np.cos(t2)], axis=2)
    # joints: pair up x-neighbours (cells (2k, y, z) and (2k + 1, y, z)): body index = x * n_side^2 + y * n_side + z
    pairs = [(x * n_side * n_side + r, (x + 1) * n_side * n_side + r) for x in range(0, n_side - 1, 2) for r in range(n_side * n_side)]
    pairs = pairs[:joints_per_world]
    joints = np.zeros((n_worlds, joints_per_world), dtype=_abi.JOINT_DTYPE)
    joints["body_a"] = -1  # unused slots are inactive
    for j, (ba, bb) in enumerate(pairs):
        joints["body_a"][:, j] = ba
        joints["body_b"][:, j] = bb
        joints["anchor_a"][:, j, :] = out["center_of_mass"][:, ba, :]
        joints["anchor_b"][:, j, :] = out["center_of_mass"][:, bb, :]
        joints["distance"][:, j] = np.linalg.norm(com_world[:, bb, :] - com_world[:, ba, :], axis=1)
    return out, hulls, np.ascontiguousarray(hull_ids), joints
