#!/usr/bin/env python
"""Per-source-line instruction / stall-sample shares from an .ncu-rep captured with --import-source on.
usage: python scripts/ncu_lines.py rep.ncu-rep [top_n]"""
import collections
import csv
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
agg, samp, thr, srcs = collections.Counter(), collections.Counter(), collections.Counter(), {}
cur, hdr = None, None
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur, hdr = r[1].split("/")[-1], None
        continue
    if r[0] == "Function Name":
        continue
    if r[0] == "Line No":
        hdr = r
        iI, iS, iT = hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Thread Instructions Executed")
        continue
    if hdr is None:
        continue
    try:
        ln, n, s, t = int(r[0]), int(r[iI]), int(r[iS]), int(r[iT])
    except ValueError:
        continue
    agg[(cur, ln)] += n; samp[(cur, ln)] += s; thr[(cur, ln)] += t
    srcs[(cur, ln)] = r[1]
tot, ts = sum(agg.values()), sum(samp.values())
print(f"total warp instructions {tot}  stall samples {ts}")
byfile, sf = collections.Counter(), collections.Counter()
for k, v in agg.items():
    byfile[k[0]] += v
for k, v in samp.items():
    sf[k[0]] += v
for f, v in byfile.most_common():
    print(f"{f:28s} {100 * v / tot:5.1f}% inst {100 * sf[f] / ts:5.1f}% samples")
print()
for k, v in agg.most_common(top):
    print("%-18s %4d  inst %5.2f%%  samp %5.2f%%  thr/inst %4.1f  %s" % (
        k[0], k[1], 100 * v / tot, 100 * samp[k] / ts, thr[k] / max(v, 1), srcs[k].strip()[:100]))
