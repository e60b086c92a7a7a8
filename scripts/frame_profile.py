"""Per-frame cost of a multi-body workload: frame by frame, the device time of the default schedule (world ranges in
flight) and, on a second batch evolving the same scene, the per-stage kernel times with each kernel alone
(xpbd_batch_set_sub_batches(1)), candidate pairs and constraints.  usage: python scripts/frame_profile.py [c3|c4] [frames]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

wl = sys.argv[1] if len(sys.argv) > 1 else "c4"
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 60
cfg = dict(bench.WORKLOADS[wl]) if hasattr(bench, "WORKLOADS") else None
if cfg is None:
    raise SystemExit("bench.WORKLOADS not found")
cfg["preroll"] = 0
torch.cuda.set_device(0)
b = bench.Bench(wl, cfg, 0, 0, 1, torch, None)
w = b.w
rows = []
for f in range(frames):
    w.set_sub_batches(0)
    state = w.download_dynamic()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(b.stream)
    w.integrate(bench.DT, b.S, 1, sync=False)
    e1.record(b.stream)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    after = w.download_dynamic()
    # the same frame again with each kernel alone (bit-identical result: the schedule does not change arithmetic)
    w.upload_dynamic(state)
    w.set_sub_batches(1)
    w.integrate(bench.DT, b.S, 1, sync=True)
    st = w.stats()
    assert np.array_equal(w.download_dynamic(), after)
    sm = st["stage_ms"]
    rows.append((f, ms, sum(sm) * b.S, sm, st["candidate_pairs"], st["constraints"]))
    print(f"frame {f:3d}  default {ms:7.3f} ms  alone {sum(sm) * b.S:7.3f} ms  per substep: prepare {sm[0]:.4f} sat {sm[1]:.4f} "
          f"manifold {sm[2]:.4f} solve {sm[3]:.4f}  candidates {st['candidate_pairs']}  constraints {st['constraints']}", flush=True)
