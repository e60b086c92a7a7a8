"""CPU-baseline timing of the oracle for bench.py (cpu_baseline leg and --impl reference) — the only
place outside tests/ and smoke() where the oracle is executed, and only as the thing compared against."""
import time

import numpy as np

import oracle_lib as xo
import oracle_scenes as xs  # scenes built with the oracle only: these legs never import or map the product library

DT = 1.0 / 60.0


class _Sample:
    """A bounded sample of a workload, prepared like the GPU arm (same scene generator, same pre-roll)."""

    def __init__(self, workload, cfg, n_worlds, dtype, threads):
        init, hulls, extras = xs.make_scene(workload, n_worlds, 0)
        self.workload, self.cfg, self.threads = workload, cfg, threads
        self.r = np.ascontiguousarray(init.reshape(-1, 38)).astype(dtype)
        self.hulls, self.extras = hulls, extras
        self.n_worlds = n_worlds
        self.step(cfg["preroll"])  # contact-active state, as on the GPU

    def step(self, frames, threads=None):
        threads = threads or self.threads
        if not frames:
            return
        if self.workload == "c2":
            xo.step_bodies(self.r, self.hulls[0].vertices, DT, self.cfg["substeps"], frames, threads=threads)
        else:
            import oracle_world as xw
            xw.step_worlds(self.r, self.cfg["bodies"], list(self.hulls), DT, self.cfg["substeps"], frames, threads=threads,
                           hull_ids=self.extras.get("hull_ids"), joints=self.extras.get("joints"))


def time_oracle(workload, cfg, seconds_target, dtype):
    """Returns (body-substeps/s, cores, sample description, extra dict)."""
    threads = xo.hardware_threads()
    B, S = cfg["bodies"], cfg["substeps"]
    n_worlds = max(threads * 8, 4096 // B) if B <= 16 else threads * 2
    out = {}
    sample = ""
    for name, dt_ in (("f64", np.float64), ("f32", np.float32)):
        smp = _Sample(workload, cfg, n_worlds, dt_, threads)
        t0 = time.perf_counter(); smp.step(1); probe = time.perf_counter() - t0
        frames = int(max(1, min(200, (seconds_target / 2) / max(probe, 1e-6))))
        t0 = time.perf_counter(); smp.step(frames); el = time.perf_counter() - t0
        out[name] = n_worlds * B * S * frames / el
        sample = (f"{n_worlds} worlds x {B} bodies x {frames} frames x {S} substeps after {cfg['preroll']} pre-roll frames, "
                  f"oracle port, {threads} threads")
    # single thread, f64
    nw1 = max(64 // B, 8) if B <= 16 else 2
    smp = _Sample(workload, cfg, nw1, np.float64, threads)
    f1 = 20 if B <= 16 else 3
    t0 = time.perf_counter(); smp.step(f1, threads=1); el = time.perf_counter() - t0
    one = nw1 * B * S * f1 / el
    return out["f64"], threads, sample, {"value_f32": out["f32"], "value_f64_1thread": one, "precision": "f64 (the reference's)"}


def run_reference_arm(workload, cfg, steps, warmup):
    """bench.py --impl reference: each step = one frame of a bounded sample, f64, all host threads."""
    threads = xo.hardware_threads()
    B, S = cfg["bodies"], cfg["substeps"]
    n_worlds = max(threads * 64, 16384 // B) if B <= 16 else threads * 4
    if workload != "c2":
        n_worlds = max(threads * 8, n_worlds // 8)  # multi-body worlds cost ~100x more per body-substep
    smp = _Sample(workload, cfg, n_worlds, np.float64, threads)
    for _ in range(warmup):
        smp.step(1)
    t0 = time.perf_counter()
    for _ in range(steps):
        smp.step(1)
    el = time.perf_counter() - t0
    return {"value": n_worlds * B * S * steps / el, "ms_per_step": el / steps * 1e3, "cores": threads,
            "sample": f"{n_worlds} worlds x {B} bodies per step (bounded sample of the workload), f64 oracle port, {threads} threads"}


def algorithmic_flops(workload, cfg, n_worlds=None):
    """Algorithmic flops of the specification per body-substep, by phase (add = mul = div = sqrt = 1, SURVEY.md section 8
    row D): the oracle instantiated with its counting scalar (oracle/xo_count.hpp) steps a small sample of the workload
    for one frame after the same pre-roll.  The reference's face / edge queries are counted as written (brute force)."""
    import oracle_world as xw
    threads = xo.hardware_threads()
    B, S = cfg["bodies"], cfg["substeps"]
    n_worlds = n_worlds or (16 if B <= 16 else 2)
    smp = _Sample(workload, cfg, n_worlds, np.float32, threads)
    c = xw.count_flops_worlds(smp.r, B, list(smp.hulls), DT, S, 1, hull_ids=smp.extras.get("hull_ids"),
                              joints=smp.extras.get("joints"), body_body=workload != "c2")
    bs = c["body_substeps"]
    per_phase = {p: sum(v.values()) / bs for p, v in c["flops"].items()}
    kinds = {k: sum(v[k] for v in c["flops"].values()) / bs for k in ("add", "mul", "div", "sqrt")}
    return {"per_body_substep": c["total"] / bs, "per_phase": per_phase, "by_kind": kinds,
            "sample": f"{n_worlds} worlds x {B} bodies x 1 frame x {S} substeps after {cfg['preroll']} pre-roll frames, oracle with counting scalar"}


def parity_sample(workload, cfg, first_world, world_ids, frames, gpu38):
    """Max position / orientation error of sampled worlds against the fp32 oracle stepped `frames` frames from the same
    scene (row E: per-GPU summary statistics).  gpu38: (len(world_ids) * B, 38) f64 array downloaded from the device."""
    import oracle_world as xw
    B, S = cfg["bodies"], cfg["substeps"]
    parts, ids, joints = [], [], []
    hulls = None
    for w in world_ids:
        init, hulls, extras = xs.make_scene(workload, 1, first_world + int(w))
        parts.append(init.reshape(-1, 38))
        if "hull_ids" in extras:
            ids.append(extras["hull_ids"]); joints.append(extras["joints"])
    r = np.ascontiguousarray(np.concatenate(parts)).astype(np.float32)
    if workload == "c2":
        xo.step_bodies(r, hulls[0].vertices, DT, S, frames, threads=xo.hardware_threads())
    else:
        xw.step_worlds(r, B, list(hulls), DT, S, frames, threads=xo.hardware_threads(),
                       hull_ids=np.concatenate(ids) if ids else None, joints=np.concatenate(joints) if joints else None)
    c = r.astype(np.float64)
    g = np.asarray(gpu38, dtype=np.float64).reshape(-1, 38)
    pos = float(np.abs(g[:, 31:34] - c[:, 31:34]).max())
    a, b = g[:, 34:38], c[:, 34:38]
    av, as_, bv, bs = a[:, :3], a[:, 3], -b[:, :3], b[:, 3]
    v = as_[:, None] * bv + bs[:, None] * av + np.cross(av, bv)
    rot = float((2.0 * np.arcsin(np.clip(np.linalg.norm(v, axis=1), 0.0, 1.0))).max())
    return pos, rot, bool(np.array_equal(g, c))
