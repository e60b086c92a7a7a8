"""CPU-baseline timing of the oracle for bench.py (cpu_baseline leg and --impl reference) — the only
place outside tests/ and smoke() where the oracle is executed, and only as the thing compared against."""
import time

import numpy as np

import oracle_lib as xo

DT = 1.0 / 60.0


def _step_sample(workload, cfg, init38, hull_verts, frames, threads):
    if workload == "c2":
        return xo.step_bodies(init38, hull_verts, DT, cfg["substeps"], frames, threads=threads)
    import oracle_world as xw
    return xw.step_worlds(init38, cfg["bodies"], hull_verts, DT, cfg["substeps"], frames, threads=threads)


def _prepare(workload, cfg, n_worlds, dtype, make_scene, threads):
    init, hulls, _ = make_scene(workload, n_worlds, 0)
    r = xo.rigid38(init).astype(dtype)
    verts = hulls[0].vertices
    # same pre-roll as the GPU arm so the timed frames are contact-active
    _step_sample(workload, cfg, r, verts, cfg["preroll"], threads)
    return r, verts


def time_oracle(workload, cfg, seconds_target, dtype, make_scene):
    """Returns (body-substeps/s, cores, sample description, extra dict)."""
    threads = xo.hardware_threads()
    B, S = cfg["bodies"], cfg["substeps"]
    n_worlds = max(threads * 8, 4096 // B)
    out = {}
    for name, dt_ in (("f64", np.float64), ("f32", np.float32)):
        r, verts = _prepare(workload, cfg, n_worlds, dt_, make_scene, threads)
        t0 = time.perf_counter(); _step_sample(workload, cfg, r, verts, 1, threads); probe = time.perf_counter() - t0
        frames = int(max(1, min(200, (seconds_target / 2) / max(probe, 1e-6))))
        t0 = time.perf_counter(); _step_sample(workload, cfg, r, verts, frames, threads); el = time.perf_counter() - t0
        out[name] = n_worlds * B * S * frames / el
        sample = f"{n_worlds} worlds x {B} bodies x {frames} frames x {S} substeps after {cfg['preroll']} pre-roll frames, oracle port, {threads} threads"
    # single thread, f64
    r, verts = _prepare(workload, cfg, max(64 // B, 8), np.float64, make_scene, threads)
    nw1 = len(r) // B
    t0 = time.perf_counter(); _step_sample(workload, cfg, r, verts, 20, 1); el = time.perf_counter() - t0
    one = nw1 * B * S * 20 / el
    return out["f64"], threads, sample, {"value_f32": out["f32"], "value_f64_1thread": one, "precision": "f64 (the reference's)"}


def run_reference_arm(workload, cfg, steps, warmup, make_scene):
    """bench.py --impl reference: each step = one frame of a bounded sample, f64, all host threads."""
    threads = xo.hardware_threads()
    B, S = cfg["bodies"], cfg["substeps"]
    n_worlds = max(threads * 64, 16384 // B)
    r, verts = _prepare(workload, cfg, n_worlds, np.float64, make_scene, threads)
    for _ in range(warmup):
        _step_sample(workload, cfg, r, verts, 1, threads)
    t0 = time.perf_counter()
    for _ in range(steps):
        _step_sample(workload, cfg, r, verts, 1, threads)
    el = time.perf_counter() - t0
    return {"value": n_worlds * B * S * steps / el, "ms_per_step": el / steps * 1e3, "cores": threads,
            "sample": f"{n_worlds} worlds x {B} bodies per step (bounded sample of the workload), f64 oracle port, {threads} threads"}
