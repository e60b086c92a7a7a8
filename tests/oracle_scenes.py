"""Synthetic scenes of BASELINE.json's configs built with the ORACLE only (test infrastructure): the same worlds as
constraint_solver_b200/scenes.py, bit for bit (tests/test_oracle_scenes.py), without importing the product — this is
what bench.py's CPU arms (`cpu_baseline`, `--impl reference`) step, so that the reference arm's process never maps
libxpbd_b200.so.  Rigids are (n_worlds, bodies, 38) f64 arrays in Rigid's field order (src/rigid.rs:9-47)."""
import json
import os

import numpy as np

import oracle_lib as xo

SEED = 0x5EED
HERE = os.path.dirname(os.path.abspath(__file__))
F = {name: slice(o, o + n) for name, o, n in xo.RIGID_FIELDS}


def cube() -> xo.Hull:
    return xo.Hull.builtin(1)


def rigid_from_hull(hull: xo.Hull, density: float) -> np.ndarray:
    """Rigid::new(polytope.rigid_metrics(density)) (src/rigid.rs:53-71, src/geometry/integrate.rs:26-75)."""
    rc, r = xo.rigid_new(xo.rigid_metrics(hull.tuple5(), density))
    assert rc == 0
    return r


def c2_perturbed(n_worlds: int, first_world: int = 0) -> np.ndarray:
    out = np.empty((n_worlds, 1, 38))
    out[:] = xo.world_new()[0]
    bitgen = np.random.Philox(key=SEED)
    bitgen.advance(first_world * 4)
    u = np.random.Generator(bitgen).random((n_worlds, 16))
    out[:, 0, 33] = 2.0 + 4.0 * u[:, 0]
    out[:, 0, F["velocity"]] = -3.0 + 6.0 * u[:, 1:4]
    out[:, 0, F["angular_velocity"]] = -5.0 + 10.0 * u[:, 4:7]
    a, b = np.sqrt(1.0 - u[:, 7]), np.sqrt(u[:, 7])
    t1, t2 = 2.0 * np.pi * u[:, 8], 2.0 * np.pi * u[:, 9]
    out[:, 0, F["rotation"]] = np.stack([a * np.sin(t1), a * np.cos(t1), b * np.sin(t2), b * np.cos(t2)], axis=1)
    return out


def c3_stack(n_worlds: int, first_world: int = 0, n_cubes: int = 10, gap: float = 1e-3, jitter: float = 0.01) -> np.ndarray:
    base = rigid_from_hull(cube(), 0.1)
    base[F["external_force"]] = (0.0, 0.0, -2.0)
    out = np.empty((n_worlds, n_cubes, 38))
    out[:] = base
    bitgen = np.random.Philox(key=SEED + 3)
    per_world = 4 * ((2 * n_cubes + 3) // 4)
    bitgen.advance(first_world * per_world // 4)
    u = np.random.Generator(bitgen).random((n_worlds, per_world))[:, : 2 * n_cubes].reshape(n_worlds, n_cubes, 2)
    out[:, :, 31:33] = -jitter + 2.0 * jitter * u
    out[:, :, 33] = gap + np.arange(n_cubes)[None, :] * (1.0 + gap)
    return out


def c4_hulls() -> list:
    """tests/golden/c4_hulls.json (written by tests/golden/make_c4_hulls.py from the product's QuickHull)."""
    with open(os.path.join(HERE, "golden", "c4_hulls.json")) as f:
        raw = json.load(f)
    fh = lambda rows: [[float.fromhex(x) for x in r] for r in rows]
    return [xo.Hull(fh(h["vertices"]), h["edges"], h["faces"], [float.fromhex(x) for x in h["centroid"]]) for h in raw]


def c4_pile(n_worlds: int, first_world: int = 0, n_side: int = 4, joints_per_world: int = 32, spacing: float = 2.2,
            jitter: float = 0.1):
    hulls = c4_hulls()
    B = n_side ** 3
    protos = []
    for k, h in enumerate(hulls):
        r = rigid_from_hull(h, 0.1 if k == 0 else 0.3)
        r[F["external_force"]] = (0.0, 0.0, -20.0 / r[0])
        protos.append(r)
    bitgen = np.random.Philox(key=SEED + 4)
    per_world = 4 * ((B * 8 + 3) // 4)
    bitgen.advance(first_world * per_world // 4)
    u = np.random.Generator(bitgen).random((n_worlds, per_world))[:, : B * 8].reshape(n_worlds, B, 8)
    hull_ids = np.minimum((u[:, :, 0] * len(hulls)).astype(np.int32), len(hulls) - 1)
    out = np.empty((n_worlds, B, 38))
    for k, p in enumerate(protos):
        out[hull_ids == k] = p
    ix, iy, iz = np.meshgrid(np.arange(n_side), np.arange(n_side), np.arange(n_side), indexing="ij")
    lattice = np.stack([ix.ravel(), iy.ravel(), iz.ravel()], axis=1).astype(np.float64)
    centre = lattice * spacing + np.array([0.0, 0.0, 1.5])
    com_world = centre[None, :, :] + (-jitter + 2.0 * jitter * u[:, :, 1:4])
    out[:, :, F["position"]] = com_world - out[:, :, F["center_of_mass"]]
    a, b = np.sqrt(1.0 - u[:, :, 4]), np.sqrt(u[:, :, 4])
    t1, t2 = 2.0 * np.pi * u[:, :, 5], 2.0 * np.pi * u[:, :, 6]
    out[:, :, F["rotation"]] = np.stack([a * np.sin(t1), a * np.cos(t1), b * np.sin(t2), b * np.cos(t2)], axis=2)
    pairs = [(x * n_side * n_side + r, (x + 1) * n_side * n_side + r) for x in range(0, n_side - 1, 2) for r in range(n_side * n_side)]
    pairs = pairs[:joints_per_world]
    joints = np.zeros((n_worlds, joints_per_world), dtype=xo.JOINT_DTYPE)
    joints["body_a"] = -1
    for j, (ba, bb) in enumerate(pairs):
        joints["body_a"][:, j] = ba
        joints["body_b"][:, j] = bb
        joints["anchor_a"][:, j, :] = out[:, ba, F["center_of_mass"]]
        joints["anchor_b"][:, j, :] = out[:, bb, F["center_of_mass"]]
        joints["distance"][:, j] = np.linalg.norm(com_world[:, bb, :] - com_world[:, ba, :], axis=1)
    return out, hulls, np.ascontiguousarray(hull_ids), joints


def make_scene(workload: str, n_worlds: int, first_world: int):
    """(rigids38 [n_worlds, B, 38] f64, hulls, extras) for bench.py's workloads c2 / c3 / c4."""
    if workload == "c2":
        return c2_perturbed(n_worlds, first_world), [cube()], {}
    if workload == "c4":
        init, hulls, ids, joints = c4_pile(n_worlds, first_world)
        return init, hulls, {"hull_ids": ids, "joints": joints}
    return c3_stack(n_worlds, first_world), [cube()], {}
