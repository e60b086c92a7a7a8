#!/usr/bin/env python
"""Experiment: does splitting the C3 batch into K sub-batches of worlds on K CUDA streams raise throughput?
(the SAT stage is issue-bound, the solve / manifold stages latency-bound: kernels of different sub-batches can share
an SM).  Prints body-substeps/s for K = 1, 2, 3, 4."""
import os, sys, time
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import constraint_solver_b200 as cs
from constraint_solver_b200 import scenes

N, B, S = int(os.environ.get("EXP_WORLDS", 100_000)), 10, 20
dt = 1.0 / 60.0
cube = cs.Polytope.new_cube()
for K in [int(k) for k in os.environ.get("EXP_K", "1,2,3,4").split(",")]:
    per = N // K
    streams = [torch.cuda.Stream() for _ in range(K)]
    ws = []
    for k in range(K):
        w = cs.Worlds(per, B, device=0, body_body_contacts=True)
        w.add_hull(cube)
        w.set_stream(streams[k].cuda_stream)
        w.upload(scenes.c3_stack(per, k * per))
        ws.append(w)
    for w in ws:
        w.integrate(dt, S, 30, sync=False)
    torch.cuda.synchronize()
    steps = 8
    for rep in range(2):
        t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        t0.record(torch.cuda.default_stream())
        for s_ in streams: s_.wait_event(t0)
        for i in range(steps):
            for w in ws:
                w.integrate(dt, S, 1, sync=False)
        for s_ in streams:
            e = torch.cuda.Event(); e.record(s_); torch.cuda.default_stream().wait_event(e)
        t1.record(torch.cuda.default_stream())
        torch.cuda.synchronize()
        ms = t0.elapsed_time(t1)
    print(f"K={K} worlds/sub-batch={per} ms/frame={ms/steps:.3f} value={per*K*B*S*steps/(ms*1e-3):.4g} "
          f"SAT_BLOCKS_PER_SM={os.environ.get('XPBD_SAT_BLOCKS_PER_SM','-')}", flush=True)
    for w in ws: w.close()
