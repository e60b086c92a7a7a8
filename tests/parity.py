"""Shared parity helpers: GPU (through the C ABI) vs the fp32 CPU oracle on the same seeded inputs."""
import numpy as np

import oracle_lib as xo

POS_TOL = 1e-4  # metres   (BASELINE.json north_star)
ROT_TOL = 1e-4  # radians


def rot_error(q_gpu: np.ndarray, q_cpu: np.ndarray) -> np.ndarray:
    """2*asin(|vec(q_gpu * conj(q_cpu))|) per row; quaternions as v.x v.y v.z s."""
    a, b = q_gpu.astype(np.float64), q_cpu.astype(np.float64)
    av, as_ = a[..., :3], a[..., 3]
    bv, bs = -b[..., :3], b[..., 3]
    v = as_[..., None] * bv + bs[..., None] * av + np.cross(av, bv)
    return 2.0 * np.arcsin(np.clip(np.linalg.norm(v, axis=-1), 0.0, 1.0))


def compare(gpu_rec: np.ndarray, cpu38_f32: np.ndarray):
    """Returns (max position error, max orientation error, all-bitwise-equal) over all bodies."""
    g = xo.rigid38(gpu_rec)
    c = cpu38_f32.astype(np.float64).reshape(-1, 38)
    pos = np.abs(g[:, 31:34] - c[:, 31:34]).max() if len(g) else 0.0
    rot = rot_error(g[:, 34:38], c[:, 34:38]).max() if len(g) else 0.0
    return float(pos), float(rot), bool(np.array_equal(g, c))


def compare_arrays(gpu38_f64: np.ndarray, cpu38_f32: np.ndarray):
    """The same for an already flattened (n, 38) GPU array."""
    g = gpu38_f64.reshape(-1, 38)
    c = cpu38_f32.astype(np.float64).reshape(-1, 38)
    pos = np.abs(g[:, 31:34] - c[:, 31:34]).max() if len(g) else 0.0
    rot = rot_error(g[:, 34:38], c[:, 34:38]).max() if len(g) else 0.0
    return float(pos), float(rot), bool(np.array_equal(g, c))
