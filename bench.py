#!/usr/bin/env python
"""bench.py — body-substeps/s of the batched XPBD step on N B200s (one process per GPU, worlds sharded by
world index, no collective on the step path), with the roofline of the step kernel and the CPU baseline.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2|c3|c4] [--impl ours|reference] [--no-extras]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
         bench.py --gpus N --steps K --warmup W

A "step" is one frame (World::integrate, src/world.rs:34-43): `substeps` XPBD substeps of every body of
every world.  Prints ONE JSON line on rank 0:
  value            device-resident throughput of the default workload C3 (CUDA events, max over ranks)
  e2e              the same through xpbd_batch_step_host_f32 (packed fp32 state in pinned host memory, H2D + D2H per step)
  e2e_rigid_f64    the same through xpbd_batch_step_host_rigid (the reference's own 304-byte f64 records, H2D + D2H per step)
  roofline         dominant kernel: FP32 issue (flops EXECUTED per candidate from the committed ncu capture x candidates per
                   launch / kernel time, against this run's non-FMA FP32 microbenchmark); HBM figures as sub-keys; the
                   algorithmic flops of the specification, counted by the oracle's counting scalar in this run, next to it
  per_rank         elapsed, body-substeps, constraints, overflow / NaN worlds, sampled parity error of every rank (all_gather)
  other_workloads  C2 and C4 measured in the same run (N = 1)
  cpu_baseline     the oracle port on the host cores, bounded sample (N = 1)
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "body_substeps_per_sec"
UNIT = "body-substeps/s"

WORKLOADS = {
    # name: worlds per GPU, bodies per world, substeps per frame, pre-roll frames (untimed, so that the timed
    # frames run on contact-active state instead of free fall)
    "c2": dict(worlds=1 << 20, bodies=1, substeps=25, preroll=45,
               desc="C2: 2^20 perturbed single-cube-on-ground worlds per GPU (BASELINE.json configs[1])"),
    "c3": dict(worlds=100_000, bodies=10, substeps=20, preroll=30,
               desc="C3/C5: 10-cube stack per world, 100k worlds per GPU, 20 substeps/frame (configs[2], [4])"),
    "c4": dict(worlds=16_384, bodies=64, substeps=20, preroll=20,
               desc="C4: 64-body pile of mixed QuickHull hulls with 32 distance joints per world, 16k worlds (configs[3])"),
}
# Default = C3: the configuration BASELINE.json's target (>= 1e9 body-substeps/s per B200) and its weak-scaling
# sweep (configs[4]: the same worlds sharded over 1/2/4/8 GPUs) are quoted on.  `--workload c2` is configs[1].
DEFAULT_WORKLOAD = "c3"
BYTES_PER_BODY_PER_STEP = 204  # 38 floats read + 13 written
PARITY_SAMPLE_WORLDS = 16      # worlds per rank compared with the fp32 oracle after the timed region (row E statistics)


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f), "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}, "fallback"


class ClockSampler(threading.Thread):
    """Samples SM clock, max SM clock and throttle reasons DURING the timed region (B200_PROFILING.md) through
    NVML (same counters as `nvidia-smi --query-gpu=clocks.sm,clocks.max.sm,clocks_event_reasons.*`), every 10 ms."""

    def __init__(self, gpu_index: int):
        super().__init__(daemon=True)
        self.gpu, self.samples, self._stop_evt = gpu_index, [], threading.Event()
        self.max_mhz, self.h, self.nv = None, None, None
        try:
            import pynvml
            pynvml.nvmlInit()
            visible = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = int(visible.split(",")[gpu_index]) if visible and visible.split(",")[0].isdigit() else gpu_index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(idx)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.nv = pynvml
        except Exception:
            self.nv = None

    def _sample(self):
        nv = self.nv
        if nv is not None:
            sm = float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
            r = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
            names = []
            for bit, name in ((nv.nvmlClocksThrottleReasonHwSlowdown, "hw_slowdown"),
                              (nv.nvmlClocksThrottleReasonHwThermalSlowdown, "hw_thermal_slowdown"),
                              (nv.nvmlClocksThrottleReasonSwThermalSlowdown, "sw_thermal_slowdown"),
                              (nv.nvmlClocksThrottleReasonSwPowerCap, "sw_power_cap")):
                if r & bit:
                    names.append(name)
            self.samples.append((sm, names))
        else:
            q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
                 "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
            out = subprocess.run(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-i", str(self.gpu)],
                                 capture_output=True, text=True, timeout=5).stdout.strip().split(",")
            out = [x.strip() for x in out]
            self.max_mhz = float(out[1])
            names = [n for n, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], out[2:6])
                     if v.lower().startswith("active")]
            self.samples.append((float(out[0]), names))

    def run(self):
        while not self._stop_evt.is_set():
            try:
                self._sample()
            except Exception:
                pass
            self._stop_evt.wait(0.01)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=6)
        sm = [s[0] for s in self.samples]
        reasons = sorted({n for s in self.samples for n in s[1]})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": self.max_mhz, "reasons": reasons,
                "samples": len(self.samples), "source": "nvml" if self.nv is not None else "nvidia-smi"}


def make_scene(workload: str, n_worlds: int, first_world: int):
    """Returns (rigids, hulls, body_body, extras) with extras = {hull_ids, joints} for C4."""
    import constraint_solver_b200 as cs
    from constraint_solver_b200 import scenes
    cube = cs.Polytope.new_cube()
    if workload == "c2":
        return scenes.c2_perturbed(n_worlds, first_world), [cube], False, {}
    if workload == "c4":
        init, hulls, ids, joints = scenes.c4_pile(n_worlds, first_world)
        return init, hulls, True, {"hull_ids": ids, "joints": joints}
    return scenes.c3_stack(n_worlds, first_world), [cube], True, {}


def _oracle_modules():
    """The CPU oracle (test infrastructure): used by bench.py only as the timed CPU baseline / reference arm, as the
    flop counter of the specification and as the checker of the per-rank parity sample — never on the measured path."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import bench_oracle
    return bench_oracle


def run_reference(args, cfg, rank):
    """--impl reference: the reference's CPU implementation of the path = the oracle port in f64 (the reference's own
    precision; the Rust crate cannot be built here), all host threads, each step one frame of a bounded sample of the same
    scene after the same pre-roll.  Scenes come from the oracle-side generators: this process never maps libxpbd_b200.so."""
    if rank != 0:
        return
    res = _oracle_modules().run_reference_arm(args.workload, cfg, args.steps, args.warmup)
    line = {
        "impl": "reference", "metric": METRIC, "value": res["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": res["ms_per_step"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(cfg, args.gpus, sample=res["sample"]),
        "cpu_baseline": {"value": res["value"], "unit": UNIT, "cores": res["cores"], "kind": "port", "sample": res["sample"]},
        "e2e": {"value": res["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


def workload_config(cfg, world_size, sample=None, sub_batches=None):
    """`config` of the JSON line: identical keys in both arms; the CPU arm differs in ONE named key, `sample`."""
    n_bodies = cfg["worlds"] * cfg["bodies"]
    c = {"workload": cfg["desc"], "worlds_per_gpu": cfg["worlds"], "bodies_per_world": cfg["bodies"], "substeps": cfg["substeps"],
         "dt": "1/60", "preroll_frames": cfg["preroll"], "parallelism": f"worlds sharded x{world_size}, no collective",
         "l2": f"inputs larger than L2: {n_bodies * 152 / 1e6:.0f} MB of body state per GPU",
         "fma": "off (bit parity with the fp32 oracle)",
         "sample": sample or "whole workload"}
    if sub_batches is not None:
        c["sub_batches"] = f"{sub_batches} contiguous world ranges per GPU stepped concurrently on {sub_batches} streams (xpbd_batch_set_sub_batches, automatic)"
    return c


_REAL_STDOUT = None


def _quiet_stdout():
    """Everything except the final JSON line goes to stderr: libraries (NCCL prints its version banner on stdout at
    communicator creation) must not pollute the one-line contract."""
    global _REAL_STDOUT
    if _REAL_STDOUT is None:
        sys.stdout.flush()
        _REAL_STDOUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)


def emit(line: dict):
    out = _REAL_STDOUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


class Bench:
    """One workload on this rank's GPU."""

    def __init__(self, workload, cfg, rank, local_rank, world_size, torch, dist):
        import constraint_solver_b200 as cs
        self.workload, self.cfg, self.rank, self.local_rank, self.world_size = workload, cfg, rank, local_rank, world_size
        self.torch, self.dist = torch, dist
        self.n_worlds, self.B, self.S = cfg["worlds"], cfg["bodies"], cfg["substeps"]
        self.n_bodies = self.n_worlds * self.B
        self.first_world = rank * self.n_worlds
        init, self.hulls, self.body_body, self.extras = make_scene(workload, self.n_worlds, self.first_world)
        joints = self.extras.get("joints")
        self.multi_body = self.body_body or joints is not None
        self.w = cs.Worlds(self.n_worlds, self.B, device=local_rank, body_body_contacts=self.body_body,
                           joints_per_world=0 if joints is None else joints.shape[1])
        for h in self.hulls:
            self.w.add_hull(h)
        if "hull_ids" in self.extras:
            self.w.set_hulls(self.extras["hull_ids"])
        if joints is not None:
            self.w.set_joints(joints)
        self.stream = torch.cuda.current_stream()
        self.w.set_stream(self.stream.cuda_stream)
        self.w.upload(init)
        self.init = init  # the f64 records a reference caller would hold (e2e_rigid_f64 uploads them every step)
        self.frames_done = 0
        if cfg["preroll"]:
            self.w.integrate(DT, self.S, cfg["preroll"])  # untimed: bring the scene into contact
            self.frames_done += cfg["preroll"]
        # every timed leg (value, e2e, e2e_rigid_f64) starts from this state and steps warmup + K frames: the legs evolve
        # the same scene over the same frames, so their numbers are comparable
        self.start_state = self.w.download_dynamic()
        self.start_frames = self.frames_done

    def rewind(self):
        self.w.upload_dynamic(self.start_state)
        self.w.sync()
        self.frames_done = self.start_frames

    def barrier(self):
        if self.world_size > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, x: float) -> float:
        t = self.torch.tensor([x], dtype=self.torch.float64, device="cuda")
        if self.world_size > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def device_resident(self, steps, warmup):
        """`value`: K step calls back to back, state resident in HBM, CUDA events on the batch stream."""
        torch, w = self.torch, self.w
        for _ in range(warmup):
            w.integrate(DT, self.S, 1, sync=False)
        self.barrier()
        sampler = ClockSampler(self.local_rank)
        sampler.start()
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
        ev[0].record(self.stream)
        for i in range(steps):
            w.integrate(DT, self.S, 1, sync=False)
            ev[i + 1].record(self.stream)
        self.barrier()
        clocks = sampler.stop()
        self.frames_done += warmup + steps
        total_ms = ev[0].elapsed_time(ev[-1])
        launch_ms = [ev[i].elapsed_time(ev[i + 1]) for i in range(steps)]
        stats = w.stats()
        total_ms_max = self.max_over_ranks(total_ms)
        value = self.world_size * self.n_bodies * self.S * steps / (total_ms_max * 1e-3)
        return {"value": value, "ms_per_step": total_ms_max / steps, "rank_ms": total_ms, "launch_ms": launch_ms, "stats": stats,
                "clocks": clocks}

    def parity(self):
        """Sampled worlds of this rank against the fp32 oracle after every frame stepped so far (row E statistics)."""
        ids = np.unique(np.linspace(0, self.n_worlds - 1, PARITY_SAMPLE_WORLDS).astype(np.int64))
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        import oracle_lib as xo
        got = np.concatenate([xo.rigid38(self.w.download(int(i), 1)) for i in ids])
        pos, rot, exact = _oracle_modules().parity_sample(self.workload, self.cfg, self.first_world, ids, self.frames_done, got)
        return {"worlds": int(len(ids)), "frames": int(self.frames_done), "max_pos_err_m": pos, "max_rot_err_rad": rot, "bit_identical": exact}

    def stage_times(self, reps=5):
        """Per-stage device times of the staged pipeline, each kernel ALONE (one chain of launches on the batch stream)."""
        if not self.multi_body:
            return None, 0
        acc, cand = np.zeros(4), 0
        self.w.set_sub_batches(1)
        for _ in range(reps):
            self.w.integrate(DT, self.S, 1, sync=True)
            st = self.w.stats()
            acc += np.array(st["stage_ms"])
            cand = int(st["candidate_pairs"])
        self.frames_done += reps
        self.w.set_sub_batches(0)
        return (acc / reps).tolist(), cand

    def e2e_f32(self, steps, warmup):
        """Per step: H2D of every body's packed fp32 dynamic state from pinned memory, one frame, D2H of the result, sync."""
        torch, w = self.torch, self.w
        self.rewind()
        pin_in = torch.empty((self.n_bodies, 13), dtype=torch.float32, pin_memory=True)
        pin_out = torch.empty((self.n_bodies, 13), dtype=torch.float32, pin_memory=True)
        w.download_dynamic_ptr(pin_in.data_ptr(), 0, self.n_worlds)
        for _ in range(warmup):
            w.integrate_host_ptr(DT, self.S, 1, pin_in.data_ptr(), pin_out.data_ptr(), sync=True)
            pin_in, pin_out = pin_out, pin_in
        self.barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            w.integrate_host_ptr(DT, self.S, 1, pin_in.data_ptr(), pin_out.data_ptr(), sync=True)
            pin_in, pin_out = pin_out, pin_in  # the state evolves exactly as in the device-resident run
        self.barrier()
        sec = self.max_over_ranks(time.perf_counter() - t0)
        self.frames_done += warmup + steps
        io = self.n_bodies * 13 * 4
        return {"value": self.world_size * self.n_bodies * self.S * steps / sec, "unit": UNIT, "h2d_bytes_per_step": io,
                "d2h_bytes_per_step": io,
                "what": "xpbd_batch_step_host_f32 (pinned host buffers: H2D of every body's dynamic state, one frame, D2H of the result) + xpbd_batch_sync per step, wall clock"}

    def e2e_rigid_f64(self, steps, warmup):
        """Per step: H2D of every body's full f64 record (the reference's `Rigid`, 304 B) from pinned memory, one frame,
        D2H of the records (13 dynamic values refreshed, f64 constants untouched), sync."""
        torch, w = self.torch, self.w
        from constraint_solver_b200 import _abi
        self.rewind()
        nbytes = self.n_bodies * _abi.RIGID_DTYPE.itemsize
        pin = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
        rec = pin.numpy().view(_abi.RIGID_DTYPE)
        rec[:] = w.download().reshape(-1)
        for _ in range(warmup):
            w.integrate_host_rigid_ptr(DT, self.S, 1, pin.data_ptr(), pin.data_ptr(), sync=True)
        self.barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            w.integrate_host_rigid_ptr(DT, self.S, 1, pin.data_ptr(), pin.data_ptr(), sync=True)  # in place, like `World` values
        self.barrier()
        sec = self.max_over_ranks(time.perf_counter() - t0)
        self.frames_done += warmup + steps
        return {"value": self.world_size * self.n_bodies * self.S * steps / sec, "unit": UNIT, "h2d_bytes_per_step": nbytes,
                "d2h_bytes_per_step": nbytes,
                "what": "xpbd_batch_step_host_rigid (the reference's own 304-byte f64 Rigid records in pinned host memory: H2D of all 38 fields, one frame, D2H of the records) + xpbd_batch_sync per step, wall clock"}

    def close(self):
        self.w.close()


def fp32_peak(local_rank):
    import ctypes
    from constraint_solver_b200 import _abi
    tf = ctypes.c_double(0.0)
    if _abi.load().xpbd_fp32_peak_microbench(local_rank, 0, ctypes.byref(tf)) == 0:
        return tf.value
    return None


def ncu_figures(workload):
    """Per-launch figures of the dominant kernel from the committed ncu capture (profiles/traffic.json)."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            return json.load(f).get(workload) or {}
    except Exception:
        return {}


def roofline(b: Bench, dev, stage_ms, cand_pairs, peak_tf, alg):
    peaks, peak_kind = measured_peaks()
    avg_launch_s = float(np.mean(dev["launch_ms"])) * 1e-3
    step_gbs = b.n_bodies * BYTES_PER_BODY_PER_STEP / avg_launch_s / 1e9
    hbm_step = {"achieved": step_gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": step_gbs / peaks["hbm_gbs"],
                "peak_source": peak_kind, "algorithmic_bytes_per_step": b.n_bodies * BYTES_PER_BODY_PER_STEP,
                "note": "204 B per body per frame (38 floats read + 13 written, DESIGN.md section 2): structurally tiny, the path is FP32-issue bound (SURVEY.md section 8 row D)"}
    ncu = ncu_figures(b.workload)
    value = dev["value"] / b.world_size  # per GPU
    r = {"bound": "fp32_issue", "unit": "TFLOP/s", "peak": peak_tf,
         "peak_source": "xpbd_fp32_peak_microbench (non-FMA FADD+FMUL chains: FMA is off on the whole path for bit parity), this run",
         "traffic": None}
    if b.multi_body and stage_ms is not None and stage_ms[1] > 0:
        kernel = ncu.get("kernel", "stage_sat_thread_kernel" if b.workload == "c3" else "stage_sat_kernel")
        exe = ncu.get("executed_flops_per_candidate")
        kernel_s = stage_ms[1] * 1e-3
        r.update({"kernel": kernel, "kernel_ms": stage_ms[1], "candidate_pairs_per_launch": cand_pairs,
                  "stage_ms": dict(zip(["prepare", "sat", "manifold", "solve"], stage_ms)),
                  "share_of_step": stage_ms[1] / max(sum(stage_ms), 1e-9),
                  "note": "dominant kernel of the staged pipeline, timed ALONE (sub-batches = 1) with CUDA events on the batch stream inside the library; achieved = flops the kernel EXECUTES per candidate (FADD + FMUL + 2 FFMA thread instructions of the committed ncu capture) x candidates of this launch / its time"})
        if exe and peak_tf:
            r["achieved"] = cand_pairs * exe / kernel_s / 1e12
            r["frac"] = r["achieved"] / peak_tf
            r["executed_flops_per_candidate"] = exe
            r["executed_flops_source"] = ncu.get("source")
        mean_verts = float(np.mean([len(h.vertices) for h in b.hulls])) if "hull_ids" not in b.extras else \
            float(np.mean(np.array([len(h.vertices) for h in b.hulls])[b.extras["hull_ids"]]))
        bytes_per_pair = 2 * 16.0 * mean_verts + 2 * 16 + 3 * 16 + 16 + 8  # float4 vertices, 2 rotations, 2 inverse frame positions + centroid, descriptor, result
        r["hbm"] = {"achieved": cand_pairs * bytes_per_pair / kernel_s / 1e9, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                    "peak_source": peak_kind, "algorithmic_bytes_per_launch": cand_pairs * bytes_per_pair, "bytes_per_pair": bytes_per_pair}
        r["hbm"]["frac"] = r["hbm"]["achieved"] / peaks["hbm_gbs"]
        if ncu.get("dram_bytes_per_launch"):
            r["traffic"] = ncu["dram_bytes_per_launch"] * (cand_pairs / max(ncu.get("candidates", cand_pairs), 1))
            r["traffic_source"] = ncu.get("source")
    else:
        kernel_s = avg_launch_s
        r.update({"kernel": "step_single_kernel", "kernel_ms": avg_launch_s * 1e3})
        if alg and peak_tf:  # the reference defines this path's flops: the oracle's count IS the algorithmic figure
            r["achieved"] = b.n_bodies * b.S * alg["per_body_substep"] / kernel_s / 1e12
            r["frac"] = r["achieved"] / peak_tf
            r["note"] = "whole step = one launch; achieved = algorithmic flops of solver::step as written (oracle counting scalar, this run) / launch time"
        if ncu.get("dram_bytes_per_launch"):
            r["traffic"] = ncu["dram_bytes_per_launch"] * (b.n_worlds / ncu["worlds"])
            r["traffic_source"] = ncu.get("source")
    r["hbm_step"] = hbm_step
    if alg:
        eq = value * alg["per_body_substep"] / 1e12
        r["specification"] = {"flops_per_body_substep": alg["per_body_substep"], "per_phase": alg["per_phase"], "by_kind": alg["by_kind"],
                              "sample": alg["sample"], "equivalent_tflops_per_gpu": eq, "equivalent_frac_of_peak": eq / peak_tf if peak_tf else None,
                              "note": "algorithmic flops of the specification AS WRITTEN (add = mul = div = sqrt = 1), counted in this run by the oracle's counting scalar; the reference's face / edge queries are brute force with every transform recomputed per use, the kernels hoist common subexpressions and cull edge pairs exactly, so the equivalent rate can exceed the FP32 peak — it says how fast the reference's arithmetic would have to run to keep up"}
    return r


def main():
    _quiet_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--workload", choices=sorted(WORKLOADS), default=os.environ.get("XPBD_BENCH_WORKLOAD", DEFAULT_WORKLOAD))
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    ap.add_argument("--worlds", type=int, default=0, help="override worlds per GPU (smaller runs for profiling)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the other workloads, the e2e legs' second form and the parity sample (profiling runs)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    cfg = dict(WORKLOADS[args.workload])
    if args.worlds:
        cfg["worlds"] = args.worlds
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world_size = int(os.environ.get("WORLD_SIZE", "1"))

    if args.impl == "reference":
        run_reference(args, cfg, rank)
        return

    import torch
    import torch.distributed as dist

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: constraint_solver_b200 has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world_size > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    b = Bench(args.workload, cfg, rank, local_rank, world_size, torch, dist)
    dev = b.device_resident(args.steps, args.warmup)
    stats = dev["stats"]
    par = None if args.no_extras else b.parity()
    stage_ms, cand_pairs = b.stage_times()
    e2e = b.e2e_f32(args.steps, args.warmup)
    e2e64 = None if args.no_extras else b.e2e_rigid_f64(args.steps, args.warmup)

    # ---- per-rank summary statistics, gathered once after the timed regions ----
    mine = [dev["rank_ms"], float(b.n_bodies * b.S * args.steps), float(stats["constraints"]), float(stats["overflow_worlds"]),
            float(stats["nan_worlds"]), par["max_pos_err_m"] if par else -1.0, par["max_rot_err_rad"] if par else -1.0,
            float(par["worlds"]) if par else 0.0, 1.0 if (par and par["bit_identical"]) else 0.0]
    t = torch.tensor(mine, dtype=torch.float64, device="cuda")
    if world_size > 1:
        allr = [torch.zeros_like(t) for _ in range(world_size)]
        dist.all_gather(allr, t)
    else:
        allr = [t]
    per_rank = [{"rank": i, "elapsed_ms": float(x[0]), "body_substeps": int(x[1]), "constraints_last_step": int(x[2]),
                 "overflow_worlds": int(x[3]), "nan_worlds": int(x[4]),
                 "parity_sample": None if x[7] == 0 else {"worlds": int(x[7]), "frames": par["frames"], "max_pos_err_m": float(x[5]),
                                                          "max_rot_err_rad": float(x[6]), "bit_identical": bool(x[8])}}
                for i, x in enumerate(a.cpu().tolist() for a in allr)]

    if rank == 0:
        peak_tf = fp32_peak(local_rank)
        alg = None
        try:
            alg = _oracle_modules().algorithmic_flops(args.workload, cfg)
        except Exception as e:  # reported, never required for the GPU number
            sys.stderr.write(f"algorithmic flop count failed: {e}\n")
        line = {
            "metric": METRIC, "value": dev["value"], "unit": UNIT, "n_gpus": world_size, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dev["ms_per_step"], "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(cfg, world_size, sub_batches=int(stats.get("sub_batches", 1))),
            "e2e": e2e, "gpu_launches": args.steps * stats["kernel_launches"], "clocks": dev["clocks"],
            "roofline": roofline(b, dev, stage_ms, cand_pairs, peak_tf, alg),
            "constraints_last_step": int(stats["constraints"]),
            "per_rank": per_rank,
        }
        if e2e64:
            line["e2e_rigid_f64"] = e2e64
    b.close()
    del b

    # ---- the other configs in the same run (N = 1): device-resident value + e2e, fewer steps ----
    if rank == 0 and world_size == 1 and not args.no_extras and not args.worlds:
        others = {}
        for wl in ("c2", "c4"):
            if wl == args.workload:
                continue
            try:
                ocfg = dict(WORKLOADS[wl])
                ob = Bench(wl, ocfg, 0, local_rank, 1, torch, dist)
                k = max(3, min(args.steps, 30))
                od = ob.device_resident(k, 3)
                oe = ob.e2e_f32(k, 3)
                ost = od["stats"]
                others[wl] = {"value": od["value"], "unit": UNIT, "ms_per_step": od["ms_per_step"], "steps": k, "warmup": 3,
                              "config": workload_config(ocfg, 1, sub_batches=int(ost.get("sub_batches", 1))),
                              "e2e": oe, "gpu_launches": k * ost["kernel_launches"], "constraints_last_step": int(ost["constraints"]),
                              "overflow_worlds": int(ost["overflow_worlds"]), "nan_worlds": int(ost["nan_worlds"]),
                              "clocks": od["clocks"]}
                ob.close()
                del ob
            except Exception as e:
                others[wl] = {"failed": str(e)}
        line["other_workloads"] = others

    if rank == 0:
        if not args.no_cpu_baseline and world_size == 1:
            try:
                v, cores, sample, extra = _oracle_modules().time_oracle(args.workload, cfg, 12.0, np.float64)
                line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample, **extra}
            except Exception as e:  # the baseline is reported, never required for the GPU number
                line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": 0, "kind": "port", "sample": f"failed: {e}"}
        emit(line)
    if world_size > 1:
        dist.destroy_process_group()


DT = 1.0 / 60.0

if __name__ == "__main__":
    main()
