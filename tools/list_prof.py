#!/usr/bin/env python
"""Per-kernel device times of one path on the bench workload (profiling aid; also the command
the ncu captures in profiles/ were taken from).

    python tools/list_prof.py [--path lists|events|dense|auto] [--execs N] [--chroms chr21,chr22]
                              [--reads N] [--max-shift D] [--opt name=value ...]
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--path", default="lists")
    ap.add_argument("--execs", type=int, default=5)
    ap.add_argument("--chroms", default="")
    ap.add_argument("--reads", type=int, default=20_000_000)
    ap.add_argument("--max-shift", type=int, default=1000)
    ap.add_argument("--mode", default="both")
    ap.add_argument("--opt", action="append", default=[])
    args = ap.parse_args()
    from pymasc_b200 import _shim, synth
    rl, D = 36, args.max_shift
    chroms = [c for c in args.chroms.split(",") if c] or None
    total = float(sum(L for _, L in synth.HG38))
    share = sum(L for n, L in synth.HG38 if chroms is None or n in chroms) / total
    wl = synth.workload(n_reads=int(args.reads * share), chroms=chroms, read_len=rl,
                        with_mappability=args.mode == "both")
    flags = _shim.FLAG_NCC | (_shim.FLAG_MSCC if args.mode == "both" else 0)
    jobs = [_shim.Job(w["length"], rl, D, flags, w["fwd"], w["rev"], w.get("begins"), w.get("ends")) for w in wl]
    dev = _shim.Device(0)
    dev.set_option("path", {"auto": 0, "dense": 1, "events": 2, "lists": 3}[args.path])
    dev.set_option("kernel_timing", 1)      # per-kernel events (off by default: they cost a few us per step)
    for kv in args.opt:
        k, v = kv.split("=")
        dev.set_option(k, int(v))
    batch = dev.prepare(jobs)
    rows = []
    for _ in range(args.execs):
        batch.execute()
        rows.append(dev.timing())
    out = batch.fetch()
    keys = ("build_ms", "ev_fwd_ms", "ev_rev_ms", "ev_edge_ms", "window_ms", "total_ms")
    best = {k: float(np.min([r[k] for r in rows[1:] or rows])) for k in keys}
    best.update(path=rows[-1]["path"], launches=rows[-1]["launches"], execs=args.execs,
                n_fwd=int(sum(j.fwd_pos.size for j in jobs)), n_rev=int(sum(j.rev_pos.size for j in jobs)),
                n_intervals=int(sum(j.map_begins.size for j in jobs)),
                ncc_sum=int(sum(int(o["ncc"].sum()) for o in out)), max_shift=D)
    print(json.dumps(best))


if __name__ == "__main__":
    main()
