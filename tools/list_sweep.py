#!/usr/bin/env python
"""Per-kernel device times of the list path over several builds of the library and option sets
(profiling aid).  The bench workload is generated once and cached as .npy files in a scratch
directory; every (library, options) pair then runs in its own process, because the library is
bound at import time (PYMASC_B200_LIB).

    python tools/list_sweep.py --libs tools/variants/lib_a.so,tools/variants/lib_b.so \
        --opts "list_blocks_per_sm=2;list_blocks_per_sm=4" [--max-shift 1000] [--reads 20000000]
        [--chroms chr1,chr2] [--share N]    (--share N: the N-th part of the genome by position, a rank's share)
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

KEYS = ("fwd", "rev", "begins", "ends")


def save_workload(directory, reads, chroms, share):
    from pymasc_b200 import synth
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import bench_blocks
    per = bench_blocks.reads_per_chrom(reads)
    idx = [i for i, (n, _) in enumerate(synth.HG38) if chroms is None or n in chroms]
    res = bench_blocks.generate([(i, per[i], 20260101, True) for i in idx])     # synth.workload's seed: the bench workload
    meta = []
    for i, _, fwd, rev, b, e in res:
        name, L = synth.HG38[i]
        for k, a in zip(KEYS, (fwd, rev, b, e)):
            np.save(os.path.join(directory, "%s_%s.npy" % (name, k)), a)
        meta.append({"name": name, "length": int(L)})
    if share > 1:       # keep the first 1/share of the concatenated genome (whole chromosomes, then a cut one)
        budget = sum(m["length"] for m in meta) // share
        for m in meta:
            m["hi"] = int(max(0, min(m["length"], budget)))
            budget -= m["hi"]
    json.dump(meta, open(os.path.join(directory, "meta.json"), "w"))


def child(directory, max_shift, execs, opts):
    from pymasc_b200 import _shim
    meta = json.load(open(os.path.join(directory, "meta.json")))
    rl = 36
    jobs = []
    for m in meta:
        a = {k: np.load(os.path.join(directory, "%s_%s.npy" % (m["name"], k))) for k in KEYS}
        hi = m.get("hi")
        if hi is not None:
            if hi <= 0:
                continue
            if hi < m["length"]:        # a rank's range: the reads and runs its halos need
                a["fwd"] = a["fwd"][a["fwd"] < hi + 1]
                a["rev"] = a["rev"][a["rev"] < hi + max_shift + 2]
                keep = a["begins"] < hi + rl + 1
                a["begins"], a["ends"] = a["begins"][keep], a["ends"][keep]
        j = _shim.Job(m["length"], rl, max_shift, _shim.FLAG_NCC | _shim.FLAG_MSCC, a["fwd"], a["rev"], a["begins"], a["ends"])
        if hi is not None and hi < m["length"]:
            j.lo, j.hi = 0, hi
        jobs.append(j)
    dev = _shim.Device(0)
    dev.set_option("path", 3)
    dev.set_option("kernel_timing", 1)      # per-kernel events (off by default; --opts kernel_timing=0 for the plain step)
    for kv in opts:
        k, v = kv.split("=")
        dev.set_option(k, int(v))
    batch = dev.prepare(jobs)
    rows = []
    for _ in range(execs):
        batch.execute()
        rows.append(dev.timing())
    out = batch.fetch()
    keys = ("ev_fwd_ms", "ev_rev_ms", "ev_edge_ms", "total_ms")
    best = {k: round(float(np.min([r[k] for r in rows[2:]])), 4) for k in keys}
    best["median_total_ms"] = round(float(np.median([r["total_ms"] for r in rows[2:]])), 4)
    h = 0
    for o in out:
        for k in ("ncc", "mscc", "mfwd", "mrev", "dmp"):
            h = (h * 1000003 + int(np.asarray(o[k], dtype=np.int64).sum())) % (1 << 61)
    best["checksum"] = h
    print(json.dumps(best))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--libs", default="")
    ap.add_argument("--opts", default="")
    ap.add_argument("--max-shift", type=int, default=1000)
    ap.add_argument("--reads", type=int, default=20_000_000)
    ap.add_argument("--chroms", default="")
    ap.add_argument("--share", type=int, default=1)
    ap.add_argument("--execs", type=int, default=12)
    ap.add_argument("--child", default="")
    args = ap.parse_args()
    if args.child:
        child(args.child, args.max_shift, args.execs, [o for o in args.opts.split(",") if o])
        return
    chroms = [c for c in args.chroms.split(",") if c] or None
    with tempfile.TemporaryDirectory(dir="/dev/shm" if os.path.isdir("/dev/shm") else None) as d:
        save_workload(d, args.reads, chroms, args.share)
        for lib in [l for l in args.libs.split(",") if l] or [""]:
            for opts in args.opts.split(";") if args.opts else [""]:
                env = dict(os.environ)
                if lib:
                    env["PYMASC_B200_LIB"] = os.path.abspath(lib)
                r = subprocess.run([sys.executable, os.path.abspath(__file__), "--child", d, "--opts", opts,
                                    "--max-shift", str(args.max_shift), "--execs", str(args.execs)],
                                   env=env, capture_output=True, text=True)
                line = r.stdout.strip().splitlines()[-1] if r.stdout.strip() else "FAILED rc=%d %s" % (r.returncode, r.stderr[-400:])
                print("%s | %s | %s" % (os.path.basename(lib) or "default", opts or "-", line), flush=True)


if __name__ == "__main__":
    main()
