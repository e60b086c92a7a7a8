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
