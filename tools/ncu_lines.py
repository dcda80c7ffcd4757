#!/usr/bin/env python
"""Per-source-line instruction and stall-sample shares from
`ncu -i X.ncu-rep --page source --csv --print-source cuda,sass --kernel-name regex:K` output
(profiling aid).  usage: ncu_lines.py FILE.csv [min_percent]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
thresh = float(sys.argv[2]) if len(sys.argv) > 2 else 0.5
hdr = None
lines = []
for r in rows:
    if r and r[0] == "Line No":
        hdr = r
        continue
    if hdr is None or len(r) < len(hdr) or not r[0]:
        continue
    try:
        ln = int(r[0])
    except ValueError:
        continue
    ie = hdr.index("Instructions Executed")
    te = hdr.index("Thread Instructions Executed")
    sm = hdr.index("# Samples")
    try:
        lines.append((ln, r[1], int(r[ie]), int(r[te]), int(r[sm])))
    except ValueError:
        pass
ti = sum(x[2] for x in lines) or 1
ts = sum(x[4] for x in lines) or 1
print("total warp instructions %d, samples %d" % (ti, ts))
for ln, src, n, t, s in lines:
    if 100.0 * n / ti >= thresh or 100.0 * s / ts >= thresh:
        print("%5.1f%% inst %5.1f%% samp  thr/inst %4.1f  L%-4d %s" %
              (100.0 * n / ti, 100.0 * s / ts, t / max(n, 1), ln, src.strip()[:100]))
