#!/usr/bin/env python
"""Per-kernel totals of an ncu launch list (`ncu --metrics gpu__time_duration.sum --csv --log-file X.csv ...`):
launches, total and mean duration, share of the list kernels' time.  usage: ncu_launches.py X.csv"""
import collections
import csv
import sys


def main():
    rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if l.startswith('"'))]
    hdr = rows[0]
    kn, mv, mu = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    tot = collections.OrderedDict()
    for r in rows[1:]:
        name = r[kn].replace("void ", "").replace("pmb::", "").split("(")[0]
        us = float(r[mv].replace(",", "")) * {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(r[mu], 1.0)
        t = tot.setdefault(name, [0, 0.0])
        t[0] += 1
        t[1] += us
    lists = sum(v[1] for k, v in tot.items() if k.startswith("list_"))
    print("# kernel, launches, total us, mean us, share of the list kernels' time")
    for k, (n, us) in sorted(tot.items(), key=lambda kv: -kv[1][1] if kv[0].startswith("list_") else 1):
        share = "%.3f" % (us / lists) if k.startswith("list_") and lists else "-"
        print("%-50s %4d %10.1f %8.2f %s" % (k, n, us, us / n, share))


if __name__ == "__main__":
    main()
