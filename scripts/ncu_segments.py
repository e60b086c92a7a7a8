#!/usr/bin/env python
"""Where a kernel's warp instructions go, from the SASS page of an .ncu-rep: consecutive SASS instructions with the same
execution count are one segment (a loop body, a branch arm).  usage: python scripts/ncu_segments.py rep.ncu-rep [min_share_pct]"""
import csv
import subprocess
import sys

rep = sys.argv[1]
min_share = float(sys.argv[2]) if len(sys.argv) > 2 else 0.4
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[1]
iI, iT, iS, iM = hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed"), hdr.index("Source"), hdr.index("# Samples")
data = [(r[iS].strip(), int(r[iI]), int(r[iT]), int(r[iM])) for r in rows[2:] if len(r) > iT]
tot, ts = sum(d[1] for d in data), sum(d[3] for d in data)
segs, cur = [], None
for i, (s, n, t, sm) in enumerate(data):
    if cur and abs(n - cur[2]) <= 0.02 * max(n, cur[2], 1):
        cur[1] = i; cur[3] += n; cur[4] += t; cur[5] += sm
    else:
        cur = [i, i, n, n, t, sm]; segs.append(cur)
print(f"# {rep}\ntotal warp instructions {tot}, stall samples {ts}, static SASS instructions {len(data)}")
for a, b, n, sn, st, sm in segs:
    if sn > min_share / 100 * tot:
        print(f"sass {a:5d}-{b:5d} ({b - a + 1:4d} instrs) executed {n:9d}x  share {100 * sn / tot:5.1f}%  samples {100 * sm / max(ts, 1):5.1f}%  lanes {st / max(sn, 1):4.1f}  first: {data[a][0][:48]}")
