#!/usr/bin/env python
"""Debug: fraction of edge pairs that survive stage 1 of the culled edge query (needs the -DXPBD_CULL_STATS build,
constraint_solver_b200/lib/debug/libxpbd_b200_stats.so; see scripts/gpu_cull_stats.sh)."""
import ctypes as C, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from constraint_solver_b200 import _abi
_abi.LIB_PATH = os.environ.get("XPBD_STATS_LIB", os.path.join(ROOT, "constraint_solver_b200", "lib", "debug", "libxpbd_b200_stats.so"))
import constraint_solver_b200 as cs
from constraint_solver_b200 import scenes
sys.path.insert(0, ROOT)
import bench

lib = _abi.load()
def stats():
    out = (C.c_ulonglong * 43)()
    lib.xpbd_debug_cull_stats(out)
    return list(out)

for wl, n in (("c3", 4096), ("c4", 512)):
    cfg = bench.WORKLOADS[wl]
    init, hulls, body_body, extras = bench.make_scene(wl, n, 0)
    joints = extras.get("joints")
    w = cs.Worlds(n, cfg["bodies"], body_body_contacts=True, joints_per_world=0 if joints is None else joints.shape[1])
    for h in hulls: w.add_hull(h)
    if "hull_ids" in extras: w.set_hulls(extras["hull_ids"])
    if joints is not None: w.set_joints(joints)
    w.upload(init)
    os.environ["XPBD_WORLD_PATH"] = "staged"
    w.integrate(1 / 60, cfg["substeps"], cfg["preroll"])
    a = stats()
    w.integrate(1 / 60, cfg["substeps"], 5)
    b = stats()
    q, kept, seen = (b[i] - a[i] for i in range(3))
    hist = [b[3 + i] - a[3 + i] for i in range(40)]
    t = [b[3 + i] - a[3 + i] for i in range(20, 29)]
    if t[0]:  # thread-per-candidate kernel (all-cube batches): class-level cull statistics, per edge query
        names = ["class tests", "class culled", "fallback: ill-conditioned class (|c|^2 <= c2min)", "fallback: well-conditioned class", "pairs kept from ill-conditioned classes",
                 "pairs kept from well-conditioned classes", "walk: axis + support scan of A", "walk: support scan of B", "edge feature chosen"]
        for nm, v in zip(names, t):
            print(f"   {nm}: {v / max(q, 1):.3f} per query")
        hist = hist[:20]
    print("   kept-per-query histogram (%):", {i: round(100.0 * h / max(q, 1), 1) for i, h in enumerate(hist) if h})
    print(f"{wl}: {q} edge queries, {seen / max(q,1):.1f} edge pairs each, {kept / max(q,1):.2f} kept by stage 1 ({100.0 * kept / max(seen,1):.1f} %)")
    w.close()
