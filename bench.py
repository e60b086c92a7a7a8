#!/usr/bin/env python
"""bench.py — body-substeps/s of the batched XPBD step on N B200s (one process per GPU, worlds sharded by
world index, no collective on the step path), with the roofline of the step kernel and the CPU baseline.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2|c3] [--impl ours|reference]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
         bench.py --gpus N --steps K --warmup W

A "step" is one frame (World::integrate, src/world.rs:34-43): `substeps` XPBD substeps of every body of
every world, fused in one kernel launch.  Prints ONE JSON line on rank 0.
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
SAT_FLOPS_PER_PAIR_EXECUTED = 9146.0      # cube-cube candidate, culled edge query (ncu, profiles/r01j_stage_sat.txt)
SAT_FLOPS_PER_PAIR_BRUTE_FORCE = 17480.0  # the same candidate through the reference's full edge-pair scan (profiles/r01h_stage_sat.txt)


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


def cpu_baseline(workload: str, cfg: dict, seconds_target: float = 12.0, dtype=np.float64):
    """The oracle port (the reference is Rust and cannot be built here) on a bounded sample of the same
    workload, on all host threads.  Returns (body-substeps/s, cores, sample description)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib as xo
    import bench_oracle
    return bench_oracle.time_oracle(workload, cfg, seconds_target, dtype)


def run_reference(args, cfg, rank):
    """--impl reference: the reference's CPU implementation of the path = the oracle port in f64 (the
    reference's own precision), all host threads, bounded sample per step."""
    if rank != 0:
        return
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import bench_oracle
    res = bench_oracle.run_reference_arm(args.workload, cfg, args.steps, args.warmup)
    line = {
        "impl": "reference", "metric": METRIC, "value": res["value"], "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": res["ms_per_step"], "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": cfg["desc"], "substeps": cfg["substeps"], "sample": res["sample"]},
        "cpu_baseline": {"value": res["value"], "unit": UNIT, "cores": res["cores"], "kind": "port", "sample": res["sample"]},
        "e2e": {"value": res["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


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
    import constraint_solver_b200 as cs

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: constraint_solver_b200 has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world_size > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    n_worlds, B, S = cfg["worlds"], cfg["bodies"], cfg["substeps"]
    n_bodies = n_worlds * B
    init, hulls, body_body, extras = make_scene(args.workload, n_worlds, rank * n_worlds)
    dt = 1.0 / 60.0

    joints = extras.get("joints")
    w = cs.Worlds(n_worlds, B, device=local_rank, body_body_contacts=body_body,
                  joints_per_world=0 if joints is None else joints.shape[1])
    for h in hulls:
        w.add_hull(h)
    if "hull_ids" in extras:
        w.set_hulls(extras["hull_ids"])
    if joints is not None:
        w.set_joints(joints)
    stream = torch.cuda.current_stream()
    w.set_stream(stream.cuda_stream)
    w.upload(init)
    del init
    if cfg["preroll"]:
        w.integrate(dt, S, cfg["preroll"])  # untimed: bring the scene into contact

    def barrier():
        if world_size > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident throughput (`value`) -------------------------------------------------
    for _ in range(args.warmup):
        w.integrate(dt, S, 1, sync=False)
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    ev[0].record(stream)
    constraints = 0
    for i in range(args.steps):
        w.integrate(dt, S, 1, sync=False)
        ev[i + 1].record(stream)
    barrier()
    clocks = sampler.stop()
    total_ms = ev[0].elapsed_time(ev[-1])
    launch_ms = [ev[i].elapsed_time(ev[i + 1]) for i in range(args.steps)]
    stats = w.stats()
    constraints = stats["constraints"]
    # per-stage device times of the staged multi-body pipeline (CUDA events on the batch stream around each kernel of
    # the last substep of a step call; xpbd_step_stats.stage_ms), averaged over a few extra untimed step calls
    stage_ms, cand_pairs = None, 0
    sub_batches = int(stats.get("sub_batches", 1))
    if body_body or joints is not None:  # the staged multi-body pipeline
        acc = np.zeros(4)
        w.set_sub_batches(1)  # each stage kernel timed ALONE: one chain of launches on the batch stream (the timed region above
        #                       runs `sub_batches` world ranges concurrently, so there a kernel's wall time includes its neighbours)
        for _ in range(5):
            w.integrate(dt, S, 1, sync=True)
            st_ = w.stats()
            acc += np.array(st_["stage_ms"])
            cand_pairs = int(st_["candidate_pairs"])
        stage_ms = (acc / 5).tolist()
        w.set_sub_batches(0)
    t = torch.tensor([total_ms], dtype=torch.float64, device="cuda")
    if world_size > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms_max = float(t.item())
    value = world_size * n_bodies * S * args.steps / (total_ms_max * 1e-3)

    # ---- end to end through the C ABI with host buffers (`e2e`) -----------------------------------
    # Per step: H2D of every body's dynamic state from pinned host memory, one frame, D2H of the result.
    pin_in = torch.empty((n_bodies, 13), dtype=torch.float32, pin_memory=True)
    pin_out = torch.empty((n_bodies, 13), dtype=torch.float32, pin_memory=True)
    w.download_dynamic_ptr(pin_in.data_ptr(), 0, n_worlds)
    for _ in range(2):
        w.integrate_host_ptr(dt, S, 1, pin_in.data_ptr(), pin_out.data_ptr(), sync=True)
        pin_in, pin_out = pin_out, pin_in
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        # xpbd_batch_step_host_f32: H2D of the state, one frame, D2H of the result (per world range, so the copies of one
        # range run under the kernels of the others), then xpbd_batch_sync: the caller owns the result before the next step
        w.integrate_host_ptr(dt, S, 1, pin_in.data_ptr(), pin_out.data_ptr(), sync=True)
        pin_in, pin_out = pin_out, pin_in
    barrier()
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
    if world_size > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = world_size * n_bodies * S * args.steps / float(t.item())
    io_bytes = n_bodies * 13 * 4

    # non-FMA FP32 issue peak of this device (FADD/FMUL chains; FMA is off on the whole path for bit parity)
    fp32_peak = None
    if rank == 0:
        import ctypes
        from constraint_solver_b200 import _abi
        tf = ctypes.c_double(0.0)
        if _abi.load().xpbd_fp32_peak_microbench(local_rank, 0, ctypes.byref(tf)) == 0:
            fp32_peak = tf.value

    if rank == 0:
        peaks, peak_kind = measured_peaks()
        avg_launch_s = float(np.mean(launch_ms)) * 1e-3
        achieved_gbs = n_bodies * BYTES_PER_BODY_PER_STEP / avg_launch_s / 1e9
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world_size, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": total_ms_max / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": cfg["desc"], "worlds_per_gpu": n_worlds, "bodies_per_world": B, "substeps": S,
                       "dt": "1/60", "preroll_frames": cfg["preroll"], "parallelism": f"worlds sharded x{world_size}, no collective",
                       "l2": f"inputs larger than L2: {n_bodies * 152 / 1e6:.0f} MB of body state per GPU",
                       "fma": "off (bit parity with the fp32 oracle)",
                       "sub_batches": f"{sub_batches} contiguous world ranges per GPU stepped concurrently on {sub_batches} streams (xpbd_batch_set_sub_batches, automatic)"},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": io_bytes, "d2h_bytes_per_step": io_bytes,
                    "what": "xpbd_batch_step_host_f32 (pinned host buffers: H2D of every body's dynamic state, one frame, D2H of the result) + xpbd_batch_sync per step, wall clock"},
            "gpu_launches": args.steps * stats["kernel_launches"],
            "clocks": clocks,
            "roofline": {"bound": "hbm", "achieved": achieved_gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                         "frac": achieved_gbs / peaks["hbm_gbs"], "traffic": None, "peak_source": peak_kind,
                         "algorithmic_bytes_per_launch": n_bodies * BYTES_PER_BODY_PER_STEP,
                         "kernel_ms": float(np.mean(launch_ms)),
                         "note": "algorithmic bytes = 204 B per body per frame (DESIGN.md section 2); the path is FP32-issue bound, not HBM bound, so this fraction is structurally small (SURVEY.md section 8 row D)"},
            "constraints_last_step": int(constraints),
        }
        # FP32 side of the roofline (north_star: "achieved fraction of the HBM/FP32 roofline").  Algorithmic flops are
        # only defined by the reference for its live single-body path: 812 + 364 k per body-substep, k = ground
        # contacts.  For the multi-body configs see the ncu counters under profiles/.
        if fp32_peak:
            fp = {"peak": fp32_peak, "unit": "TFLOP/s", "peak_source": "xpbd_fp32_peak_microbench (non-FMA FADD+FMUL chains), this run"}
            if args.workload == "c2":
                flops = 812.0 * n_bodies * S + 364.0 * float(constraints)
                fp["achieved"] = flops / avg_launch_s / 1e12
                fp["frac"] = fp["achieved"] / fp32_peak
                fp["algorithmic_flops_per_launch"] = flops
            line["roofline"]["fp32"] = fp
        if stage_ms is not None and stage_ms[1] > 0:
            # Dominant kernel of the multi-body pipeline: the SAT stage — stage_sat_cubes_kernel (two cube-cube candidates per
            # warp) when every hull of the batch is cube-shaped (C3/C5), else stage_sat_kernel (one warp per candidate pair).
            sat_kernel = "stage_sat_cubes_kernel" if args.workload == "c3" else "stage_sat_kernel"
            # Algorithmic bytes per pair: world-space vertices of both hulls (12 B each), two rotations (16 B each), two
            # inverse-frame positions and the centroid of A (12 B each), the pair record (16 B read + 8 B written).
            mean_verts = float(np.mean([len(h.vertices) for h in hulls])) if "hull_ids" not in extras else \
                float(np.mean(np.array([len(h.vertices) for h in hulls])[extras["hull_ids"]]))
            bytes_per_pair = 2 * 12.0 * mean_verts + 2 * 16 + 3 * 12 + 24
            sat_gbs = cand_pairs * bytes_per_pair / (stage_ms[1] * 1e-3) / 1e9
            line["roofline"] = {
                "bound": "hbm", "kernel": sat_kernel, "achieved": sat_gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                "frac": sat_gbs / peaks["hbm_gbs"], "traffic": None, "peak_source": peak_kind,
                "algorithmic_bytes_per_launch": cand_pairs * bytes_per_pair, "kernel_ms": stage_ms[1],
                "candidate_pairs_per_launch": cand_pairs, "bytes_per_pair": bytes_per_pair,
                "stage_ms": dict(zip(["prepare", "sat", "manifold", "solve"], stage_ms)),
                "share_of_step": stage_ms[1] / max(sum(stage_ms), 1e-9),
                "step": line["roofline"],
                "note": "dominant kernel of the staged pipeline, timed ALONE (sub-batches = 1) with CUDA events on the batch stream inside the library; "
                        "issue bound (C3: 372 FADD/FMUL/FFMA of 880 warp-instructions per pair, DESIGN.md section 5), so the HBM fraction is small by construction",
            }
            if "fp32" in line["roofline"]["step"]:
                line["roofline"]["fp32"] = line["roofline"]["step"].pop("fp32")
                if args.workload == "c3":
                    # FP32 work of one cube-cube candidate, thread-level flops (FADD + FMUL + 2 x FFMA) from the ncu SASS page of
                    # stage_sat_cubes_kernel: 9 146 executed with the witness cull (profiles/r01j_stage_sat.txt), 17 480 for
                    # the reference's brute-force face + edge queries (profiles/r01h_stage_sat.txt); DESIGN.md section 5.
                    fp = line["roofline"]["fp32"]
                    fp["flops_per_pair"] = SAT_FLOPS_PER_PAIR_EXECUTED
                    fp["achieved"] = cand_pairs * SAT_FLOPS_PER_PAIR_EXECUTED / (stage_ms[1] * 1e-3) / 1e12
                    fp["frac"] = fp["achieved"] / fp["peak"]
                    fp["kernel"] = sat_kernel
                    fp["brute_force_flops_per_pair"] = SAT_FLOPS_PER_PAIR_BRUTE_FORCE
                    fp["brute_force_equivalent"] = cand_pairs * SAT_FLOPS_PER_PAIR_BRUTE_FORCE / (stage_ms[1] * 1e-3) / 1e12
                    fp["note"] = ("executed flops: the witness cull drops ~89 % of the edge pairs before the reference arithmetic; "
                                  "brute_force_equivalent = the rate at which the reference's full scan would have to run to finish in the same time")
        # roofline.traffic: DRAM bytes per launch of the dominant kernel from the committed ncu capture (profiles/)
        try:
            with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
                tr = json.load(f).get(args.workload)
            if tr:
                line["roofline"]["traffic"] = tr["dram_bytes_per_launch"] * (n_worlds / tr["worlds"])
                line["roofline"]["traffic_source"] = tr["source"]
        except Exception:
            pass
        if not args.no_cpu_baseline and world_size == 1:
            try:
                v, cores, sample, extra = cpu_baseline(args.workload, cfg)
                line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample, **extra}
            except Exception as e:  # the baseline is reported, never required for the GPU number
                line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": 0, "kind": "port", "sample": f"failed: {e}"}
        emit(line)
    w.close()
    if world_size > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
