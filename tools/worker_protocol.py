#!/usr/bin/env python
"""The `-p N` shape a maintainer gets after the one-line swap of INTEGRATION.md: N spawned worker
processes, each with its own calculator (and therefore its own pmb_ctx / CUDA context), pull
chromosomes from an order queue and answer `flush(chrom)` + `get_result(chrom)` per chromosome
through a report queue -- PyMaSC/handler/worker.py:68-104, :217-234 and handler/calc.py:163-235,
with pysam's read loop replaced by `feed_packed` of pre-decoded arrays (pysam is absent; the feed
is not what is measured here).  One device batch of ONE job per chromosome, 24 per genome.

    python tools/worker_protocol.py [--workers 1,4,8] [--gpus 1]      -> one JSON line

Reported per N: wall time of the whole genome (first order to last result), the workers' own
time in flush + get_result (steady state, and the first call of each worker, which pays library
load + CUDA context creation), next to ONE batch of all 24 chromosomes in one process.
"""
import argparse
import json
import multiprocessing as mp
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

READ_LEN, MAX_SHIFT = 36, 1000


class Track(object):
    """bwfeeder of a worker: the chromosome's intervals arrive with the order."""

    def __init__(self):
        self.tracks = {}

    def fetch_arrays(self, thr, chrom):
        if chrom not in self.tracks:
            raise KeyError(chrom)
        return self.tracks[chrom]

    def close(self):
        pass


class Quiet(object):
    def __getattr__(self, name):
        return lambda *a, **k: None


def worker(index, n_gpus, references, lengths, order_queue, report_queue):
    """CalcWorker.run (handler/worker.py:68-104): calculator per process, chromosome per order."""
    os.environ["PYMASC_B200_DEVICE"] = str(index % max(1, n_gpus))     # worker_index % n_gpus
    if ROOT not in sys.path:
        sys.path.insert(0, ROOT)
    from pymasc_b200 import CCBitArrayCalculator
    track = Track()
    calc = CCBitArrayCalculator(MAX_SHIFT, READ_LEN, references, lengths, track, False, None, Quiet())
    first = True
    while True:
        order = order_queue.get()
        if order is None:
            break
        chrom, fwd, rev, b, e = order
        track.tracks[chrom] = (b, e)
        t0 = time.perf_counter()
        calc.feed_packed(chrom, fwd, READ_LEN * fwd.size, rev, READ_LEN * rev.size,
                         int(min(fwd[0], rev[0] - READ_LEN + 1)), int(max(fwd[-1], rev[-1] - READ_LEN + 1)))
        t1 = time.perf_counter()
        calc.flush(chrom)                       # _report_result, handler/worker.py:224-234
        result = calc.get_result(chrom)
        t2 = time.perf_counter()
        report_queue.put((chrom, result, index, first, t1 - t0, t2 - t1))
        first = False
    report_queue.put(None)


def run_workers(n_workers, n_gpus, workload):
    ctx = mp.get_context("spawn")
    order_queue, report_queue = ctx.Queue(), ctx.Queue()
    references = [w["name"] for w in workload]
    lengths = [w["length"] for w in workload]
    procs = [ctx.Process(target=worker, args=(i, n_gpus, references, lengths, order_queue, report_queue))
             for i in range(n_workers)]
    for p in procs:
        p.start()
    time.sleep(3.0)                             # interpreters up (imports are not part of the measurement)
    t0 = time.perf_counter()
    for w in sorted(workload, key=lambda w: -w["length"]):      # handler/calc.py: longest first
        order_queue.put((w["name"], w["fwd"], w["rev"], w["begins"], w["ends"]))
    for _ in procs:
        order_queue.put(None)
    results, rows, done = {}, [], 0
    while done < n_workers:
        item = report_queue.get()
        if item is None:
            done += 1
            continue
        chrom, result, index, first, feed_s, flush_s = item
        results[chrom] = result
        rows.append((index, first, feed_s, flush_s))
        if len(results) == len(workload):
            wall = time.perf_counter() - t0
    for p in procs:
        p.join()
    steady = [r[3] for r in rows if not r[1]]
    firsts = [r[3] for r in rows if r[1]]
    return {"workers": n_workers, "gpus": n_gpus, "genome_wall_s": wall,
            "flush_get_result_ms_steady_median": 1e3 * float(np.median(steady)) if steady else None,
            "flush_get_result_ms_steady_max": 1e3 * float(np.max(steady)) if steady else None,
            "flush_get_result_ms_steady_sum": 1e3 * float(np.sum(steady)),
            "first_flush_per_worker_s_median": float(np.median(firsts)),
            "feed_packed_ms_sum": 1e3 * float(sum(r[2] for r in rows)),
            "chromosomes": len(results)}, results


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workers", default="1,4,8")
    ap.add_argument("--gpus", type=int, default=1)
    args = ap.parse_args()
    import bench
    from pymasc_b200 import CCBitArrayCalculator, _shim
    workload, _ = bench.build_workload("both")
    out = {"workload": "hg38-shaped genome, 20 M reads, NCC+MSCC, max_shift 1000 (bench.py's)", "runs": []}
    # ONE batch of all 24 chromosomes in one process, through the same calculator
    track = Track()
    for w in workload:
        track.tracks[w["name"]] = (w["begins"], w["ends"])
    for rep in range(3):
        calc = CCBitArrayCalculator(MAX_SHIFT, READ_LEN, [w["name"] for w in workload],
                                    [w["length"] for w in workload], track, False, None, Quiet())
        for w in workload:
            calc.feed_packed(w["name"], w["fwd"], READ_LEN * w["fwd"].size, w["rev"], READ_LEN * w["rev"].size,
                             1, int(w["length"]))
        t0 = time.perf_counter()
        calc.finishup_calculation()
        whole = calc.get_whole_result()
        single = time.perf_counter() - t0
    out["single_process_one_batch_ms"] = 1e3 * single
    ref = {c: r for c, r in whole.chroms.items()}
    for n in [int(x) for x in args.workers.split(",")]:
        row, results = run_workers(n, args.gpus, workload)
        row["ccbins_equal_single_batch"] = all(list(results[c].chrom.ccbins) == list(ref[c].ccbins) for c in ref)
        out["runs"].append(row)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
