"""The parts of bench.py's JSON line beyond the headline workload: BASELINE.json configs[3]
(deep library, position ranges) and configs[4] (16 read sets over one mappability track, by
chromosome) at the N the bench was launched with, and the "s per BAM" half of the metric.

Everything here is measurement scaffolding around the product's public API (pymasc_b200._shim
through the C ABI, pymasc_b200.pipeline for files); the oracle is only ever the checker of a
bounded sample, after the timed regions.
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

READ_LEN = 36


# ---------------------------------------------------------------------------- synthetic inputs
def _gen_chrom(task):
    """One chromosome of a synthetic workload (runs in a spawned worker)."""
    i, L, n, seed, with_track = task
    if ROOT not in sys.path:
        sys.path.insert(0, ROOT)
    from pymasc_b200 import synth
    fwd, rev = synth.chip_reads(i, L, n, READ_LEN, seed=seed)
    b = e = None
    if with_track:
        b, e = synth.mappability(i, L)
    return i, seed, fwd, rev, b, e


def generate(tasks, procs=None):
    """[(chrom_index, seed, fwd, rev, begins, ends)] for tasks (chrom_index, n_reads, seed,
    with_track), generated on `procs` spawned workers (numpy sorts of tens of millions of reads
    are the slow part of every block here; nothing of this is timed)."""
    import multiprocessing as mp
    from pymasc_b200 import synth
    lengths = [L for _, L in synth.HG38]
    full = [(i, lengths[i], n, seed, wt) for i, n, seed, wt in tasks]
    if not full:
        return []
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    procs = max(1, min(procs or cores, len(full)))
    if procs == 1:
        return [_gen_chrom(t) for t in full]
    order = sorted(range(len(full)), key=lambda k: -full[k][2])       # longest first
    with mp.get_context("spawn").Pool(procs) as pool:
        res = pool.map(_gen_chrom, [full[k] for k in order], chunksize=1)
    out = [None] * len(full)
    for k, r in zip(order, res):
        out[k] = r
    return out


def reads_per_chrom(genome_reads):
    from pymasc_b200 import synth
    total = float(sum(L for _, L in synth.HG38))
    return [max(8, int(round(genome_reads * L / total))) for _, L in synth.HG38]


def _empty(name, L, with_track=True):
    z = np.zeros(0, np.uint32)
    return {"name": name, "length": L, "fwd": z, "rev": z,
            "begins": z if with_track else None, "ends": z if with_track else None}


# ---------------------------------------------------------------------------- timing one job list
def time_jobs(dev, stream, jobs, steps, warmup, reduce=None, barrier=None, max_over_ranks=None,
              h2d_groups=4):
    """Resident ms/step (CUDA events on the launching stream) and end-to-end ms/step (pinned host
    arrays -> prepare / execute / [reduce] / fetch) of one rank's job list; both as the max over
    ranks when the callbacks are given."""
    import torch
    import bench
    barrier = barrier or torch.cuda.synchronize
    mx = max_over_ranks or (lambda x: x)
    h2d = bench.pinned_pack(jobs)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        batch = dev.prepare(jobs)

        def step():
            batch.execute()
            if reduce:
                reduce(batch)

        for _ in range(max(1, warmup)):
            step()
        barrier()
        ev0.record(stream)
        for _ in range(steps):
            step()
        ev1.record(stream)
        barrier()
        resident_ms = mx(ev0.elapsed_time(ev1)) / steps
        timing = dev.timing()
        outs = batch.fetch()
        batch.free()

        dev.set_option("async_h2d", 1 if h2d_groups > 0 else 0)
        if h2d_groups > 0:
            dev.set_option("h2d_groups", h2d_groups)

        def e2e_step():
            b = dev.prepare(jobs)
            try:
                b.execute()
                if reduce:
                    reduce(b)
                return b.fetch()
            finally:
                b.free()

        e2e_step()
        barrier()
        t0 = time.perf_counter()
        ev0.record(stream)
        for _ in range(steps):
            e2e_out = e2e_step()
        ev1.record(stream)
        barrier()
        wall_ms = (time.perf_counter() - t0) * 1e3
        e2e_ms = mx(max(wall_ms, ev0.elapsed_time(ev1))) / steps
        dev.set_option("async_h2d", 0)
    same = all(np.array_equal(np.asarray(a[k]), np.asarray(b[k])) for a, b in zip(outs, e2e_out) for k in a)
    d2h = sum(sum(v.size * 8 for v in o.values() if isinstance(v, np.ndarray)) + 16 for o in outs)
    return {"resident_ms": resident_ms, "e2e_ms": e2e_ms, "outs": outs, "timing": timing,
            "h2d_bytes": int(h2d), "d2h_bytes": int(d2h), "resident_equals_e2e": bool(same)}


# ---------------------------------------------------------------------------- oracle on a sample
def oracle_sample_equal(dev, chrom, read_len, max_shift, flags, lo, hi):
    """One more device job -- `chrom` restricted to the counted range [lo, hi) -- against the
    CPU oracle on the same range, every counter at every shift (after the timed regions; the
    oracle is the checker, never the thing measured)."""
    from oracle import cc_oracle
    from pymasc_b200 import _shim
    mscc = bool(flags & _shim.FLAG_MSCC)
    got = dev.run_batch([_shim.Job(chrom["length"], read_len, max_shift, flags, chrom["fwd"], chrom["rev"],
                                   chrom["begins"] if mscc else None, chrom["ends"] if mscc else None, lo, hi)])[0]
    want = cc_oracle.chrom_counts(chrom["length"], read_len, max_shift, chrom["fwd"], chrom["rev"],
                                  chrom["begins"] if mscc else None, chrom["ends"] if mscc else None,
                                  skip_ncc=not flags & _shim.FLAG_NCC, lo=lo, hi=hi)
    return sorted(got) == sorted(want) and all(np.array_equal(np.asarray(got[k]), np.asarray(want[k])) for k in want)


# ---------------------------------------------------------------------------- configs[3]
def deep_library_block(dev, stream, world, rank, comm, steps=3, warmup=2, genome_reads=200_000_000,
                       max_shift=5000, sample_bp=1_500_000):
    """BASELINE.json configs[3]: 200 M reads, NCC+MSCC, max_shift 5000, the genome cut into
    `world` position ranges with max_shift halos (pymasc_b200.sharding.plan: equal cuts of the
    concatenated bit arrays; the synthetic genome is uniform, so they are also balanced in reads),
    per-shift counters summed over the ranks.  A rank generates only the chromosomes its range
    touches.  `comm`: dict of callbacks reduce(batch) / barrier() / max_over_ranks(x) / all_true(b)."""
    from pymasc_b200 import _shim, sharding, synth
    t0 = time.time()
    lengths = [L for _, L in synth.HG38]
    names = [n for n, _ in synth.HG38]
    pieces = sharding.plan(lengths, READ_LEN, max_shift, world)[rank]
    mine = sorted({ci for ci, _, _ in pieces})
    per = reads_per_chrom(genome_reads)
    gen = {r[0]: r for r in generate([(ci, per[ci], 20260101, True) for ci in mine])}
    workload = []
    for ci, (name, L) in enumerate(synth.HG38):
        if ci in gen:
            _, _, fwd, rev, b, e = gen[ci]
            workload.append({"name": name, "length": L, "fwd": fwd, "rev": rev, "begins": b, "ends": e})
        else:
            workload.append(_empty(name, L))
    jobs = sharding.jobs_for_rank(workload, pieces, READ_LEN, max_shift)
    gen_s = time.time() - t0
    evals = synth.evals(lengths, READ_LEN, max_shift)
    r = time_jobs(dev, stream, jobs, steps, warmup, comm["reduce"] if world > 1 else None,
                  comm["barrier"], comm["max_over_ranks"])
    ci, lo, hi = pieces[0]
    ok = oracle_sample_equal(dev, workload[ci], READ_LEN, max_shift, 3, lo, min(hi, lo + sample_bp))
    n_pos = sum(j.fwd_pos.size + j.rev_pos.size for j in jobs)
    return {
        "config": "BASELINE configs[3]: deep library (200 M reads), NCC+MSCC, max_shift %d, %d position "
                  "range(s) with max_shift halos%s" % (max_shift, world, " + all-reduce of the counters" if world > 1 else ""),
        "n_gpus": world, "evals_per_step": evals, "steps": steps, "warmup": warmup,
        "ms_per_step": r["resident_ms"], "value": evals / (r["resident_ms"] * 1e-3), "unit": "evals/s",
        "e2e": {"ms_per_step": r["e2e_ms"], "value": evals / (r["e2e_ms"] * 1e-3),
                "h2d_bytes_per_step_rank0": r["h2d_bytes"], "d2h_bytes_per_step_rank0": r["d2h_bytes"]},
        "path": {1: "dense", 2: "events", 3: "lists"}.get(r["timing"]["path"]),
        "launches_per_step": r["timing"]["launches"],
        "rank0_positions": int(n_pos), "resident_equals_e2e": r["resident_equals_e2e"],
        "counters_equal_oracle_on_sample": bool(comm["all_true"](ok)),
        "oracle_sample": "every rank: the first %d bp of its own range as one more device job vs cc_oracle, "
                         "all five vectors and both totals at all %d shifts" % (sample_bp, max_shift + 1),
        "synthetic_generation_s": round(gen_s, 1), "chromosomes_generated_rank0": [names[c] for c in mine],
    }


# ---------------------------------------------------------------------------- configs[4]
def multi_file_block(dev, stream, world, rank, comm, steps=3, warmup=2, files=16, max_shift=1000,
                     sample_bp=1_500_000):
    """BASELINE.json configs[4]: 16 read sets (seeds 20260101 + 100 f) sharing one mappability
    track, MSCC only (skip_ncc), max_shift 1000, whole chromosomes per GPU (longest-first packing,
    the reference's -p N unit): a chromosome's 16 jobs form one PMB_FLAG_SAME_TRACK run, so its
    track is built (or, on the list path, just read) once.  No collective: results are per
    chromosome and stay with their owner."""
    from pymasc_b200 import _shim, sharding, synth
    t0 = time.time()
    lengths = [L for _, L in synth.HG38]
    names = [n for n, _ in synth.HG38]
    pieces = sharding.plan_by_chromosome(lengths, READ_LEN, max_shift, world)[rank]
    mine = [ci for ci, _, _ in pieces]
    per = reads_per_chrom(20_000_000)
    tasks = [(ci, per[ci], 20260101 + 100 * f, f == 0) for ci in mine for f in range(files)]
    gen = {(r[0], r[1]): r for r in generate(tasks)}
    S = _shim.FLAG_SAME_TRACK
    jobs = []
    sample = None
    for ci in mine:
        _, _, _, _, b, e = gen[(ci, 20260101)]
        for f in range(files):
            _, _, fwd, rev, _, _ = gen[(ci, 20260101 + 100 * f)]
            jobs.append(_shim.Job(lengths[ci], READ_LEN, max_shift, _shim.FLAG_MSCC | (S if f else 0), fwd, rev,
                                  None if f else b, None if f else e))
            if sample is None and f == min(3, files - 1):
                sample = {"length": lengths[ci], "fwd": fwd, "rev": rev, "begins": b, "ends": e}
    gen_s = time.time() - t0
    evals = synth.evals(lengths, READ_LEN, max_shift) * files
    r = time_jobs(dev, stream, jobs, steps, warmup, None, comm["barrier"], comm["max_over_ranks"])
    ok = True
    if sample is not None:
        lo = sample["length"] // 3 // 128 * 128
        ok = oracle_sample_equal(dev, sample, READ_LEN, max_shift, _shim.FLAG_MSCC, lo, lo + sample_bp)
    return {
        "config": "BASELINE configs[4]: %d read sets sharing one mappability track, MSCC only, max_shift %d, "
                  "whole chromosomes on %d GPU(s), no collective" % (files, max_shift, world),
        "n_gpus": world, "files": files, "evals_per_step": evals, "steps": steps, "warmup": warmup,
        "ms_per_step": r["resident_ms"], "ms_per_file": r["resident_ms"] / files,
        "value": evals / (r["resident_ms"] * 1e-3), "unit": "evals/s",
        "e2e": {"ms_per_step": r["e2e_ms"], "value": evals / (r["e2e_ms"] * 1e-3),
                "h2d_bytes_per_step_rank0": r["h2d_bytes"], "d2h_bytes_per_step_rank0": r["d2h_bytes"]},
        "path": {1: "dense", 2: "events", 3: "lists"}.get(r["timing"]["path"]),
        "launches_per_step": r["timing"]["launches"], "jobs_rank0": len(jobs),
        "resident_equals_e2e": r["resident_equals_e2e"],
        "counters_equal_oracle_on_sample": bool(comm["all_true"](ok)),
        "oracle_sample": "every rank: read set 3 of its first chromosome, %d bp, as one more device job vs "
                         "cc_oracle (dmp, mfwd, mrev, mscc, totals at all %d shifts)" % (sample_bp, max_shift + 1),
        "synthetic_generation_s": round(gen_s, 1), "chromosomes_rank0": [names[c] for c in mine],
    }


# ---------------------------------------------------------------------------- s per BAM
def s_per_bam_block(workload, max_shift=1000, directory=None, repeat=2, threads=0):
    """The second half of BASELINE.json's metric: "s per BAM vs CPU -p N".  The bench workload
    is written to disk as a coordinate-sorted BAM (BGZF) and a BigWig (bench scaffolding:
    tools/bam_e2e.py's writers), then timed from the FILES to the genome-wide result objects
    through pymasc_b200.pipeline.run_files -- native BGZF / BigWig decode, read-length estimate,
    feed rules, ONE device batch -- the reference's per-file workflow (handler/calc.py:60-161).
    Returns the seconds per stage; bench.py puts the reference's correlation stage beside it."""
    import shutil
    import tempfile
    import bam_e2e
    from pymasc_b200 import pipeline, synth
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    threads = threads or cores
    own = directory is None
    directory = directory or tempfile.mkdtemp(prefix="pmb_bench_")
    try:
        t0 = time.time()
        bam = os.path.join(directory, "synth.bam")
        bw = os.path.join(directory, "synth_36mer.bigwig")
        n_reads = bam_e2e.write_bam(bam, workload, threads)
        n_iv = bam_e2e.write_bigwig(bw, workload)
        gen_s = time.time() - t0
        runs = []
        for _ in range(repeat):
            t0 = time.perf_counter()
            results, info = pipeline.run_files([bam], bw, max_shift=max_shift, mapq_criteria=1, threads=threads)
            info["seconds"]["wall"] = time.perf_counter() - t0
            runs.append(info["seconds"])
        best = min(runs, key=lambda s: s["wall"])
        whole = results[0]
        return {
            "metric": "s per BAM (files on disk -> genome-wide NCC+MSCC result objects)", "unit": "s",
            "value": best["wall"], "higher_is_better": False, "stages_s": best, "runs": len(runs),
            "reads_in_file": int(n_reads), "intervals": int(n_iv), "max_shift": max_shift,
            "bam_bytes": os.path.getsize(bam), "bigwig_bytes": os.path.getsize(bw), "host_threads": threads,
            "read_len_estimated": info["files"][0]["read_len"],
            "result": {"type": type(whole).__name__, "forward_sum": int(whole.forward_sum),
                       "reverse_sum": int(whole.reverse_sum), "chroms": len(whole.chroms)},
            "file_generation_s": round(gen_s, 1),
            "note": "synthetic BAM: constant names and qualities, so it compresses 18:1 and inflates faster "
                    "than a real file would",
        }
    finally:
        if own:
            shutil.rmtree(directory, ignore_errors=True)
