#!/usr/bin/env python
"""bench.py -- the correlation stage of PyMaSC's bit-array path on B200s (and the reference
CPU arm beside it).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--mode both|ncc]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (BASELINE.json configs[2]/[1]): synthetic hg38-shaped genome -- 24
chromosomes with the real hg38 lengths, 20 M single-end 36 bp ChIP-like reads, synthetic
36-mer mappability track (~79 % mappable, ~1.6 M intervals), max_shift 1000, NCC+MSCC
(``--mode ncc``: NCC only, configs[1]).  One STEP = the whole correlation stage of that genome:
from the packed position / interval lists to every per-shift counter of every chromosome, and for
N > 1 the all-reduce of the counters.  N ranks cut the genome into N position ranges
(pymasc_b200/sharding.py), so total work is fixed: strong scaling.

Three device algorithms give the same counters (include/pymasc_b200.h, "path"); `auto` -- what
the product runs -- picks the list kernels for this workload.  The line says which path ran
(config.path) and carries the other two, run on the same inputs, under `paths` with
`counters_equal_*` (a full-size cross-check of three independent algorithms).

metric   bit-position x shift evals/s: sum_chrom (L + read_len + max_shift + 100) x (max_shift+1)
         evals per step / step time (delivered evals: the counts are those of the reference's
         bit-array loop, whatever the algorithm).
value    inputs (packed uint32 positions / intervals) resident in HBM; device time by CUDA
         events on the launching stream, max over ranks.
e2e      same step through the C ABI from pinned HOST arrays: H2D + kernels (+ all-reduce) +
         D2H of the counters inside the timed region.  PCIe-bound (90 MB per step).
roofline of the dominant kernel of the path that ran.  List path: `bound` hbm -- SURVEY 8(d)'s
         algorithmic bytes (the bit arrays the reference streams, 3 x Lbits/8) over the forward-read
         kernel's CUDA-event time against MEASURED_PEAKS.json, with the DRAM traffic measured by
         ncu (profiles/r02_traffic_lists.json) and the shared-memory-atomic rate that really binds
         it beside it.  paths.dense.roofline is the north-star figure: the shift-window kernel
         against the POPC rate measured live (pmb_pipe_probe).
cpu_baseline / --impl reference: the UNMODIFIED reference calculator compiled into oracle/_ref
         (CCBitArrayCalculator, fed read by read as handler/read.py does) on ALL of chr21 of the
         workload -- not a cache-resident prefix; the N = 1 line also runs that chromosome through
         the CUDA path and compares every vector (gpu_matches_reference_on_sample).
configs  BASELINE.json configs[3] (200 M reads, max_shift 5000, position ranges) and configs[4]
         (16 read sets over one track, by chromosome) at THIS N, each with ms_per_step, e2e and an
         oracle check on a sample (tools/bench_blocks.py); skipped past --budget-s.
s_per_bam  (N = 1) the workload as BAM + BigWig files on disk -> result objects, seconds per stage.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

T_START = time.time()
READ_LEN = 36
MAX_SHIFT = 1000
N_READS = 20_000_000
METRIC = "bit-position x shift evals/s"
UNIT = "evals/s"


# ---------------------------------------------------------------------------- helpers
def env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


class ClockSampler(object):
    """SM clock and throttle reasons DURING the timed region, polled through NVML every few
    milliseconds on a host thread (the timed region is only tens of ms long, shorter than
    nvidia-smi's sampling period)."""
    REASONS = ((0x8, "hw_slowdown"), (0x40, "hw_thermal_slowdown"), (0x20, "sw_thermal_slowdown"),
               (0x4, "sw_power_cap"), (0x80, "hw_power_brake_slowdown"))

    def __init__(self, index):
        self.index, self.samples, self.stop_flag, self.thread, self.err = index, [], False, None, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_sm = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception as e:      # noqa: BLE001
            self.nv, self.err = None, repr(e)

    def _sample(self, power=False):
        nv = self.nv
        try:
            sm = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
            rs = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
            pw = nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0 if power else 0.0
            self.samples.append((sm, rs, pw))
            return True
        except Exception as e:  # noqa: BLE001
            self.err = repr(e)
            return False

    def _poll(self):
        # NVML queries take a driver lock that delays work submission on the device (measured:
        # a 2 ms poll added 0.45 ms to a 2.4 ms step), so poll sparsely: one sample when the
        # region starts, one every PERIOD seconds, one (with power) when it ends.
        while not self.stop_flag:
            if not self._sample():
                return
            time.sleep(self.PERIOD)

    PERIOD = 0.02

    def start(self):
        if self.nv is not None:
            self.thread = threading.Thread(target=self._poll, daemon=True)
            self.thread.start()
        return self

    def stop(self):
        self.stop_flag = True
        if self.nv is not None:
            self._sample(power=True)     # the device is still busy: the timed work has not drained yet
        if self.thread is not None:
            self.thread.join(timeout=2)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvml unavailable: %s" % self.err]}
        sm = [x[0] for x in self.samples]
        mask = 0
        for x in self.samples:
            mask |= x[1]
        return {"sm_mhz": float(np.median(sm)), "sm_min_mhz": float(min(sm)), "sm_max_mhz": float(self.max_sm),
                "power_w_max": max(x[2] for x in self.samples), "samples": len(sm),
                "reasons": [name for bit, name in self.REASONS if mask & bit],
                "source": "nvml, polled (start / every 20 ms / end) over the same steps repeated back to back right "
                          "after the timed ones, as many as fill 0.25 s: queries inside the timed region perturb it"}


def pinned_u32(a):
    """Copy of `a` in page-locked host memory, as a numpy uint32 view (torch owns the pages)."""
    import torch
    a = np.ascontiguousarray(a, dtype=np.uint32)
    t = torch.empty(max(a.size, 1), dtype=torch.int32).pin_memory()
    v = t.numpy().view(np.uint32)[:a.size]
    v[:] = a
    v_base = v
    _PINNED.append(t)
    return v_base


_PINNED = []


def pinned_pack(jobs):
    """All position / interval arrays of the jobs in ONE page-locked buffer, back to back with
    16-byte padding (fwd, rev, begins, ends per job): the C ABI then moves a whole group of
    chromosomes with one copy.  Returns the payload bytes (padding not counted)."""
    import torch
    names = ("fwd_pos", "rev_pos", "map_begins", "map_ends")
    total = sum((getattr(j, n).size + 3) & ~3 for j in jobs for n in names if getattr(j, n) is not None)
    t = torch.empty(max(total, 4), dtype=torch.int32).pin_memory()
    buf = t.numpy().view(np.uint32)
    _PINNED.append(t)
    off = payload = 0
    for j in jobs:
        for n in names:
            a = getattr(j, n)
            if a is None:
                continue
            v = buf[off:off + a.size]
            v[:] = a
            setattr(j, n, v)
            off += (a.size + 3) & ~3
            payload += a.size * 4
    return payload


def build_workload(mode, chrom_indices=None):
    from pymasc_b200 import synth
    names = [n for n, _ in synth.HG38]
    chroms = None if chrom_indices is None else [names[i] for i in sorted(chrom_indices)]
    t0 = time.time()
    # read counts are allotted against the WHOLE genome, whichever chromosomes are generated
    total = float(sum(L for _, L in synth.HG38))
    out = []
    for i, (name, L) in enumerate(synth.HG38):
        if chroms is not None and name not in chroms:
            out.append({"name": name, "length": L, "fwd": np.zeros(0, np.uint32),
                        "rev": np.zeros(0, np.uint32),
                        "begins": np.zeros(0, np.uint32) if mode == "both" else None,
                        "ends": np.zeros(0, np.uint32) if mode == "both" else None})
            continue
        n = max(8, int(round(N_READS * L / total)))
        fwd, rev = synth.chip_reads(i, L, n, READ_LEN)
        b = e = None
        if mode == "both":
            b, e = synth.mappability(i, L)
        out.append({"name": name, "length": L, "fwd": fwd, "rev": rev, "begins": b, "ends": e})
    return out, time.time() - t0


# ---------------------------------------------------------------------------- CPU arms
REF_KEYS = ("ncc", "mscc", "mfwd", "mrev", "mappable_len", "fwd_total", "rev_total")


def _ref_task(args):
    """One worker task of the reference arm: a chromosome through the unmodified reference
    calculator (feed read by read, then finishup) -- runs in a spawned process like a CalcWorker
    (handler/worker.py:68-104).  Returns (evals, seconds, every result vector of the chromosome)."""
    (name, L, fwd, rev, begins, ends, mode, max_shift) = args
    if ROOT not in sys.path:
        sys.path.insert(0, ROOT)
    from oracle import ref_runner
    pos, ln, is_rev = ref_runner.merged_feed_order(
        fwd, np.full(len(fwd), READ_LEN), np.asarray(rev, np.int64) - READ_LEN + 1,
        np.full(len(rev), READ_LEN))
    tracks = {name: (begins, ends)} if mode == "both" else None
    t0 = time.perf_counter()
    # popcnt=None: the build is chosen per worker process through PYMASC_REF_POPCNT
    calc = ref_runner.run_reference(max_shift, READ_LEN, [name], [L], {name: (pos, ln, is_rev)},
                                    tracks=tracks, popcnt=None)
    dt = time.perf_counter() - t0
    res = calc.get_result(name)
    evals = (L + READ_LEN + max_shift + 100) * (max_shift + 1)
    vec = {"ncc": [int(x) for x in res.chrom.ccbins], "fwd_total": int(res.chrom.forward_sum),
           "rev_total": int(res.chrom.reverse_sum)}
    if mode == "both":
        m = res.mappable_chrom
        vec.update(mscc=[int(x) for x in m.ccbins], mfwd=[int(x) for x in m.forward_sum],
                   mrev=[int(x) for x in m.reverse_sum], mappable_len=[None if x is None else int(x) for x in m.mappable_len])
    return evals, dt, vec


def device_vectors(dev, task):
    """The same chromosome through the CUDA path, in the reference's result layout."""
    from pymasc_b200 import _shim
    (name, L, fwd, rev, begins, ends, mode, max_shift) = task
    flags = _shim.FLAG_NCC | (_shim.FLAG_MSCC if mode == "both" else 0)
    got = dev.run_batch([_shim.Job(L, READ_LEN, max_shift, flags, fwd, rev, begins, ends)])[0]
    vec = {"ncc": got["ncc"].tolist(), "fwd_total": int(got["fwd_total"]), "rev_total": int(got["rev_total"])}
    if mode == "both":
        d = got["dmp"].tolist()       # popc(D_k) -> the reference's mappable_len list (mscc.pyx:271, :292-298)
        n = min(READ_LEN, max_shift + 1)
        ml = [None] * READ_LEN
        ml[READ_LEN - n:READ_LEN] = d[:n][::-1]
        vec.update(mscc=got["mscc"].tolist(), mfwd=got["mfwd"].tolist(), mrev=got["mrev"].tolist(),
                   mappable_len=ml + d[2 * READ_LEN - 1:max_shift + 1])
    return vec


def whole_chromosome(chrom):
    return (chrom["name"], int(chrom["length"]), chrom["fwd"], chrom["rev"], chrom["begins"], chrom["ends"])


def sample_chromosome(chrom, length):
    """The leading `length` bp of a workload chromosome as a stand-alone chromosome."""
    fwd = chrom["fwd"][chrom["fwd"] < length]
    rev = chrom["rev"][chrom["rev"] < length + READ_LEN - 1]
    b = e = None
    if chrom["begins"] is not None:
        keep = chrom["ends"] <= length
        b, e = chrom["begins"][keep], chrom["ends"][keep]
    return (chrom["name"], int(length), fwd, rev, b, e)


def cpu_baseline_leg(workload, mode, check_device=None, with_popcnt=True):
    """Reference calculator on ONE core over ALL of chr21 of this workload at the workload's own
    max_shift (46.7 Mbp: 5.8 MB per bit array, several arrays per shift -- not cache-resident;
    BASELINE.md: "time chr1 and chr21 in full").  Rank 0, N = 1 only, after every timed region."""
    from oracle import ref_runner
    if not ref_runner.available():
        return {"value": None, "unit": UNIT, "cores": 0, "kind": "reference",
                "sample": "oracle/_ref missing (build it where /root/reference exists)"}
    chrom = workload[20]   # chr21
    task = whole_chromosome(chrom) + (mode, MAX_SHIFT)
    evals, dt, vec = _ref_task(task)
    out = {"value": evals / dt, "unit": UNIT, "cores": 1, "kind": "reference",
           "seconds": round(dt, 2), "same_config_as_gpu_arm": True,
           "sample": "ALL of %s of this workload (%d bp, %d fwd + %d rev reads, %s intervals), all %d "
                     "shifts, feed_*_read per read + finishup_calculation, %s; the genome is %.1fx this" %
                     (task[0], task[1], len(task[2]), len(task[3]),
                      "no" if task[4] is None else len(task[4]), MAX_SHIFT + 1,
                      "NCC+MSCC" if mode == "both" else "NCC",
                      sum(c["length"] for c in workload) / float(task[1]))}
    # the same chromosome through the reference rebuilt with -mpopcnt (its setup.py does
    # not pass it, so the stock build popcounts in software, external/bitcnt.h:26-44), in a
    # spawned interpreter because one process can import only one build of the package
    if with_popcnt and ref_runner.available(popcnt=True):
        import multiprocessing as mp
        os.environ["PYMASC_REF_POPCNT"] = "1"
        try:
            with mp.get_context("spawn").Pool(1) as pool:
                ev2, dt2, vec2 = pool.apply(_ref_task, (task,))
        finally:
            os.environ.pop("PYMASC_REF_POPCNT", None)
        out["reference_plus_popcnt"] = {"value": ev2 / dt2, "unit": UNIT, "cores": 1, "seconds": round(dt2, 2),
                                        "flags": "-O3 -ffast-math -mpopcnt (oracle/_ref_popcnt)",
                                        "same_counts": vec2 == vec}
    if check_device is not None:   # the same chromosome through the CUDA path must agree -- everywhere
        got = device_vectors(check_device, task)
        keys = [k for k in REF_KEYS if k in vec]
        out["gpu_matches_reference_on_sample"] = all(got[k] == vec[k] for k in keys)
        out["gpu_vs_reference_compared"] = "%s: every value of %s at all %d shifts" % (task[0], ", ".join(keys), MAX_SHIFT + 1)
    return out


def reference_arm(args):
    """--impl reference: the reference's own `-p N` scheme (spawned workers pulling one
    chromosome each, handler/calc.py:163-192) on all host cores.  Each task is ALL of chr21 of the
    workload (arrays of 5.8 MB: not cache-resident); what is bounded, so that K + W steps end
    within a few minutes, is the number of SHIFTS per step -- the reference's cost per shift does
    not depend on max_shift (ten whole-array passes per shift, mscc.pyx:288-317), so evals/s carry
    over.  With a budget that allows it the step runs the workload's own max_shift."""
    import multiprocessing as mp
    from oracle import ref_runner
    mode = args.mode
    if not ref_runner.available():
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref (compiled reference) missing"}))
        return
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    if args.ref_cores:
        cores = args.ref_cores
    workload, _ = build_workload(mode, chrom_indices=[20])
    chr21 = whole_chromosome(workload[20])
    ctx = mp.get_context("spawn")
    times = []
    with ctx.Pool(processes=cores) as pool:
        # import warm-up + calibration under full load: every core runs all of chr21 at 16 and at
        # 64 shifts; the difference is 48 shifts without the set-up (allocation, feed, track load)
        t0 = time.perf_counter()
        pool.map(_ref_task, [chr21 + (mode, 15)] * cores, chunksize=1)
        t1 = time.perf_counter()
        pool.map(_ref_task, [chr21 + (mode, 63)] * cores, chunksize=1)
        t2 = time.perf_counter()
        per_shift = max(1e-4, ((t2 - t1) - (t1 - t0)) / 48.0)
        setup = max(0.0, (t1 - t0) - 16 * per_shift)
        budget = min(40.0, 200.0 / max(1, args.steps)) - setup
        shifts = int(max(32, min(MAX_SHIFT + 1, budget / per_shift)))
        D = shifts - 1
        tasks = [chr21 + (mode, D)] * cores
        warm = [chr21 + (mode, 15)] * cores
        for s in range(args.warmup):
            pool.map(_ref_task, warm, chunksize=1)
        for s in range(args.steps):
            t0 = time.perf_counter()
            res = pool.map(_ref_task, tasks, chunksize=1)
            times.append((sum(r[0] for r in res), time.perf_counter() - t0))
    evals = sum(t[0] for t in times)
    secs = sum(t[1] for t in times)
    value = evals / secs
    sample = ("%d tasks per step (one per core, spawned workers), each ALL of chr21 of the workload (%d bp, "
              "%d fwd + %d rev reads%s) at %d of the %d shifts (max_shift %d; the reference's cost per shift "
              "is constant), feed per read + finishup, %s; warm-up steps run 16 shifts" %
              (cores, chr21[1], len(chr21[2]), len(chr21[3]),
               ", %d intervals" % len(chr21[4]) if chr21[4] is not None else "", shifts, MAX_SHIFT + 1, D,
               "NCC+MSCC" if mode == "both" else "NCC"))
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * secs / max(1, len(times)),
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u64",
        "data": "synthetic", "config": workload_config(mode),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "reference",
                         "sample": sample, "shifts_per_step": shifts, "full_max_shift": shifts == MAX_SHIFT + 1},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "whole_genome_correlation_stage_s_at_this_rate": sum(
            (L + READ_LEN + MAX_SHIFT + 100) for L in genome_lengths()) * (MAX_SHIFT + 1) / value,
    }
    print(json.dumps(line))


def genome_lengths():
    from pymasc_b200 import synth
    return [L for _, L in synth.HG38]


def workload_config(mode):
    return {"workload": "synthetic hg38-shaped genome (24 chromosomes, 3.09 Gb), 20M single-end 36 bp "
                        "reads%s, max_shift 1000, %s" %
                        (", synthetic 36-mer mappability track" if mode == "both" else "",
                         "NCC+MSCC" if mode == "both" else "NCC only"),
            "read_len": READ_LEN, "max_shift": MAX_SHIFT, "n_reads": N_READS,
            "mode": "ncc+mscc" if mode == "both" else "ncc",
            "l2": "inputs larger than L2: 5 bit arrays x 386 MB per step vs 126 MB L2; no explicit flush"}


# ---------------------------------------------------------------------------- GPU arm
def gpu_arm(args):
    import torch
    import torch.distributed as dist
    from pymasc_b200 import _shim, sharding, synth

    world = env_int("WORLD_SIZE", 1)
    rank = env_int("RANK", 0)
    local = env_int("LOCAL_RANK", 0)
    sys.stdout.flush()
    real_stdout = os.dup(1)      # libraries (NCCL banner) print to fd 1: keep it for the JSON line only
    os.dup2(2, 1)
    if world != args.gpus and world > 1:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d" % (args.gpus, world))
    if args.gpus > 1 and world == 1:
        raise SystemExit("launch with torch.distributed.run --nproc-per-node %d" % args.gpus)
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    mode = args.mode

    lengths = [L for _, L in synth.HG38]
    # every rank generates the whole (seeded) genome: the position plan weighs reads and run
    # edges as well as length, so all ranks need the same view of the data to agree on the cuts
    workload, gen_s = build_workload(mode)
    pieces = sharding.plan_weighted(workload, READ_LEN, MAX_SHIFT, world)[rank]
    jobs = sharding.jobs_for_rank(workload, pieces, READ_LEN, MAX_SHIFT)
    # page-locked copies of the packed inputs: what the host side hands to the C ABI
    h2d = pinned_pack(jobs)
    evals_total = synth.evals(lengths, READ_LEN, MAX_SHIFT)
    P = 5 if mode == "both" else 1

    dev = _shim.Device(local)
    stream = torch.cuda.Stream(device=local)
    dev.set_stream(stream.cuda_stream)
    if args.variant is not None:
        dev.set_option("variant", args.variant)
    if args.tile_words:
        dev.set_option("tile_words", args.tile_words)
    dev.set_option("path", {"auto": 0, "dense": 1, "events": 2, "lists": 3}[args.path])
    for kv in args.opt:
        k, v = kv.split("=")
        dev.set_option(k, int(v))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    with torch.cuda.stream(stream):
        # ---- resident-input arm
        if world > 1:
            # the C ABI's own communicator (pmb_comm_init): rank 0's id travels through torch.distributed,
            # which this script otherwise uses for the barrier and the max-over-ranks of the timings only
            idt = torch.zeros(_shim.Device.COMM_ID_BYTES, dtype=torch.uint8, device="cuda")
            if rank == 0:
                idt.copy_(torch.frombuffer(bytearray(_shim.Device.comm_get_id()), dtype=torch.uint8))
            dist.broadcast(idt, src=0)
            dev.comm_init(idt.cpu().numpy().tobytes(), rank, world)

        def reduce(b):
            """pmb_allreduce: in-place NCCL sum of the batch's device counter block over the ranks,
            on the context's stream.  The first reduce of a batch also settles the ranks' status
            words (range errors, the list path's order check); later ones are asynchronous."""
            b.allreduce()

        batch = dev.prepare(jobs)

        def step():
            batch.execute()
            if world > 1:
                reduce(batch)

        for _ in range(args.warmup):
            step()
        barrier()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record(stream)
        for _ in range(args.steps):
            step()
        ev1.record(stream)
        # Clocks: NVML queries take a driver lock that stalls submission / NCCL progress (measured:
        # two queries inside the region added 0.45 ms to every 2.4 ms step at N=2), so they are
        # taken over the SAME steps run again back to back, the device never going idle between --
        # and, K steps of this workload being a few milliseconds, over as many of them as fill
        # a quarter of a second (the same count on every rank: it derives from the max over ranks).
        sampler = ClockSampler(local).start() if (rank == 0 and not args.no_clocks) else None
        for _ in range(args.steps):
            step()
        barrier()
        dev_ms = max_over_ranks(ev0.elapsed_time(ev1))
        ms_per_step = dev_ms / args.steps
        clock_steps = args.steps
        if not args.no_clocks:
            more = int(min(4000, max(0.0, 250.0 / max(ms_per_step, 1e-3) - args.steps)))
            for _ in range(more):
                step()
            clock_steps += more
            if sampler is not None:     # the steps are only enqueued: keep sampling while the device works through them
                ev_c = torch.cuda.Event()
                ev_c.record(stream)
                t_c = time.perf_counter()
                while not ev_c.query() and time.perf_counter() - t_c < 1.0:
                    time.sleep(0.005)
        clocks = sampler.stop() if sampler else None      # (the last sample: right behind the last step)
        if clocks is not None:
            clocks["steps_sampled_over"] = clock_steps
        barrier()
        launches_per_step = dev.timing()["launches"]

        # ---- per-kernel times of the same step (library CUDA events on the same stream)
        win_ms, build_ms, fwd_ms, rev_ms, edge_ms = [], [], [], [], []
        dev.set_option("kernel_timing", 1)      # (list path: events between the kernels, off in the timed steps above)
        for _ in range(max(3, min(args.steps, 10))):
            batch.execute()
            t = dev.timing()
            win_ms.append(t["window_ms"]); build_ms.append(t["build_ms"])
            fwd_ms.append(t["ev_fwd_ms"]); rev_ms.append(t["ev_rev_ms"]); edge_ms.append(t["ev_edge_ms"])
        dev.set_option("kernel_timing", 0)
        pairs = t["window_word_shift_pairs"]
        win_launches = t["window_launches"]
        used_path = t["path"]
        # N > 1: the same K steps without the all-reduce (slowest rank) -> what the collective costs
        exec_only_ms = None
        if world > 1:
            barrier()
            ev0.record(stream)
            for _ in range(args.steps):
                batch.execute()
            ev1.record(stream)
            barrier()
            mine_ms = ev0.elapsed_time(ev1) / args.steps
            exec_only_ms = {"max": max_over_ranks(mine_ms), "min": -max_over_ranks(-mine_ms)}
            reduce(batch)               # the last execute above left this rank's partial counts
        resident = batch.fetch()
        batch.free()

        # ---- the other algorithms on the same inputs: identical counters are required; the dense
        #      shift-window kernel is the one the popc-pipe roofline (north_star) is quoted on
        others = {}
        if not args.single_path:
            for op in (1, 2, 3):
                if op == used_path:
                    continue
                dev.set_option("path", op)
                ob = dev.prepare(jobs)
                o_win, o_build, o_fwd = [], [], []
                for _ in range(4):
                    ob.execute()
                    t2 = dev.timing()
                    o_win.append(t2["window_ms"]); o_build.append(t2["build_ms"]); o_fwd.append(t2["ev_fwd_ms"])
                if world > 1:
                    reduce(ob)
                o_out = ob.fetch()
                ob.free()
                others[op] = {"window_ms": float(np.mean(o_win[1:])), "build_ms": float(np.mean(o_build[1:])),
                              "fwd_ms": float(np.mean(o_fwd[1:])), "pairs": t2["window_word_shift_pairs"],
                              "launches": t2["launches"],
                              "same": all(np.array_equal(np.asarray(a[k]), np.asarray(b2[k]))
                                          for a, b2 in zip(resident, o_out) for k in a)}
            dev.set_option("path", {"auto": 0, "dense": 1, "events": 2, "lists": 3}[args.path])

        # ---- end-to-end arm: pinned host arrays in, counters on the host out
        def e2e_step():
            b = dev.prepare(jobs)                      # H2D
            try:
                b.execute()
                if world > 1:
                    reduce(b)
                return b.fetch()                       # D2H + sync
            finally:
                b.free()

        # the input copy is pipelined against the kernels in groups of chromosomes
        # (what pmb_run_batch does by itself); --h2d-groups 0 = copy first, then compute
        if args.h2d_groups > 0:
            dev.set_option("async_h2d", 1)
            dev.set_option("h2d_groups", args.h2d_groups)
        for _ in range(min(args.warmup, 3)):
            e2e_out = e2e_step()
        barrier()
        t0 = time.perf_counter()
        ev0.record(stream)
        for _ in range(args.steps):
            e2e_out = e2e_step()
        ev1.record(stream)
        barrier()
        wall_ms = (time.perf_counter() - t0) * 1e3
        e2e_ms = max_over_ranks(max(wall_ms, ev0.elapsed_time(ev1))) / args.steps
        d2h = sum(sum(v.size * 8 for v in o.values() if isinstance(v, np.ndarray)) + 16 for o in e2e_out)
        dev.set_option("async_h2d", 0)

        # resident and e2e arms must agree with each other (full-size consistency check)
        same = all(np.array_equal(np.asarray(a[k]), np.asarray(b[k]))
                   for a, b in zip(resident, e2e_out) for k in a)

        popc_peak, _ = dev.pipe_probe(0)
        lop3_peak, _ = dev.pipe_probe(1)

    line = None
    if rank == 0:
        win_s = float(np.mean(win_ms)) * 1e-3
        # this rank's share of the algorithmic work (ranks are balanced to 128 bits)
        evals_rank = evals_total / world
        popc_alg = evals_rank * P / 32.0
        arrays = 5 if mode == "both" else 2
        alg_bytes = arrays * sum(L + READ_LEN + MAX_SHIFT + 100 for L in lengths) / 8.0 / world
        peaks = {}
        try:
            with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as fh:
                peaks = json.load(fh)
        except OSError:
            pass
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        peak_src = "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "6650 GB/s (of fallback)"
        bits_bytes = sum(L + READ_LEN + MAX_SHIFT + 100 for L in lengths) / 8.0 / world   # one bit array

        def dense_roofline(window_s, word_pairs):
            """Shift-window kernel: integer AND+popcount, bound by the POPC/ALU pipes."""
            return {"bound": "popc", "kernel": "shift_window_kernel",
                    "achieved": popc_alg / window_s / 1e9, "peak": popc_peak / 1e9, "unit": "Gpopc32/s",
                    "frac": popc_alg / window_s / popc_peak,
                    "peak_source": "pmb_pipe_probe(POPC) measured in this run (MEASURED_PEAKS.json has no "
                                   "integer-pipe figure); LOP3 rate %.0f Gop/s. frac > 1 is possible: "
                                   "carry-save counters execute 0.5 POPC per algorithmic popc32"
                                   % (lop3_peak / 1e9),
                    "algorithmic_popc_per_eval": P / 32.0, "kernel_ms": window_s * 1e3,
                    "executed_word_shift_pairs": int(word_pairs),
                    "traffic": 1.954e9 / world if mode == "both" else None,
                    "traffic_source": "profiles/r01_ncu_shift_window_dense_v1.txt (ncu --set full, N=1)",
                    "hbm": {"algorithmic_bytes": alg_bytes, "achieved_gbs": alg_bytes / window_s / 1e9,
                            "peak_gbs": hbm_peak, "frac": alg_bytes / window_s / 1e9 / hbm_peak}}

        def events_roofline(fwd_s, all_s):
            """Event-window kernels: every array is streamed once, windows come from L1/L2; the
            floor is HBM.  Dominant kernel: the forward-read kernel (F, R [, M, Mr])."""
            # F, R and M are each streamed from HBM once (M's tile and M's window region are the
            # same bytes: the second read hits L2)
            nbytes = (3 if mode == "both" else 2) * bits_bytes
            return {"bound": "hbm", "kernel": "event_window_kernel<EV_FWD>",
                    "achieved": nbytes / fwd_s / 1e9, "peak": hbm_peak, "unit": "GB/s",
                    "frac": nbytes / fwd_s / 1e9 / hbm_peak, "peak_source": peak_src,
                    "algorithmic_bytes": nbytes, "kernel_ms": fwd_s * 1e3,
                    "all_event_kernels_ms": all_s * 1e3,
                    "traffic": 1.168e9 if (world == 1 and mode == "both") else None,
                    "traffic_source": "profiles/r01_ncu_event_window_v20.txt (ncu --set full, N=1: "
                                      "dram__bytes_read.sum + dram__bytes_write.sum of the FWD launch)",
                    "note": "latency/issue-bound today (see profiles/); evals/s of this path are "
                            "delivered evals: the counts equal the dense kernel's bit for bit"}

        def measured_traffic():
            """DRAM bytes of every kernel of one list-path step, from the committed ncu capture
            (profiles/r02_traffic_lists.json, written by tools/ncu_traffic.py from an
            `ncu --set full` report of tools/list_prof.py) -- measured, per launch, N = 1."""
            try:
                with open(os.path.join(ROOT, "profiles", "r02_traffic_lists.json")) as fh:
                    return json.load(fh)
            except (OSError, ValueError):
                return None

        def lists_roofline(fwd_s, all_s):
            """List kernels: no bit array exists.  Per SURVEY 8(d) the algorithmic bytes of a step
            are the bit arrays the reference's path streams (2 or 3 x Lbits/8): `achieved` is those
            bytes over the dominant kernel's time, i.e. DELIVERED bandwidth (the counts equal the
            dense kernel's bit for bit).  What the kernels really read are the packed lists (90 MB);
            they are bound by shared-memory atomics and dependent issue, not by HBM -- see `lists`."""
            nbytes = (3 if mode == "both" else 2) * bits_bytes
            tr = measured_traffic() if (world == 1 and mode == "both") else None
            out = {"bound": "hbm", "kernel": "list_fwd_kernel",
                   "achieved": nbytes / fwd_s / 1e9, "peak": hbm_peak, "unit": "GB/s",
                   "frac": nbytes / fwd_s / 1e9 / hbm_peak, "peak_source": peak_src,
                   "algorithmic_bytes": nbytes, "kernel_ms": fwd_s * 1e3, "all_list_kernels_ms": all_s * 1e3,
                   "whole_step_frac": nbytes / (ms_per_step * 1e-3) / 1e9 / hbm_peak,
                   "traffic": tr["kernels"]["list_fwd_kernel"]["dram_bytes"] if tr else None,
                   "traffic_whole_step": tr["step_dram_bytes"] if tr else None,
                   "traffic_source": tr["source"] if tr else None,
                   "lists": {"input_bytes": int(h2d), "input_gbs_over_all_kernels": h2d / all_s / 1e9,
                             "note": "delivered evals: the path reads the sorted lists, never the bit arrays; "
                                     "DRAM traffic is far BELOW the algorithmic bytes"}}
            if tr and "shared_atomic_instr" in tr["kernels"]["list_fwd_kernel"]:
                n_at = tr["kernels"]["list_fwd_kernel"]["shared_atomic_instr"]
                peak_at = tr.get("atoms_warp_instr_per_s_peak")
                out["lists"].update(fwd_shared_atomic_warp_instr=n_at, fwd_atomic_instr_per_s=n_at / fwd_s,
                                    atomic_peak_instr_per_s=peak_at,
                                    atomic_frac=(n_at / fwd_s / peak_at) if peak_at else None,
                                    atomic_peak_source=tr.get("atoms_peak_source"))
            return out

        names = {1: "dense", 2: "events", 3: "lists"}
        mine = {"window_ms": win_s * 1e3, "build_ms": float(np.mean(build_ms)),
                "launches_per_step": int(launches_per_step)}
        if used_path == 3:
            roof = lists_roofline(float(np.mean(fwd_ms)) * 1e-3, win_s)
            roof.update(rev_kernel_ms=float(np.mean(rev_ms)), edge_and_finalize_ms=float(np.mean(edge_ms)),
                        zero_ms=float(np.mean(build_ms)), kernel_launches_per_step=int(launches_per_step))
        elif used_path == 2:
            roof = events_roofline(float(np.mean(fwd_ms)) * 1e-3, win_s)
            roof.update(rev_kernel_ms=float(np.mean(rev_ms)), edge_kernel_ms=float(np.mean(edge_ms)),
                        build_ms=float(np.mean(build_ms)), kernel_launches_per_step=int(win_launches))
        else:
            roof = dense_roofline(win_s, pairs)
            roof.update(build_ms=float(np.mean(build_ms)), kernel_launches_per_step=int(win_launches))
        paths = {names[used_path]: mine}
        for op, o in others.items():
            entry = {"window_ms": o["window_ms"], "build_ms": o["build_ms"], "launches_per_step": o["launches"],
                     "counters_equal_%s_path" % names[used_path]: o["same"]}
            if op == 1:     # north_star: the dense kernel's popc-pipe roofline is always reported
                entry["roofline"] = dense_roofline(o["window_ms"] * 1e-3, o["pairs"])
            if op == 2:
                entry["roofline"] = events_roofline(o["fwd_ms"] * 1e-3, o["window_ms"] * 1e-3)
            paths[names[op]] = entry
        line = {
            "metric": METRIC, "value": evals_total / (ms_per_step * 1e-3), "unit": UNIT,
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "u32",
            "data": "synthetic", "config": dict(workload_config(mode),
                                                parallelism="position ranges x%d + NCCL all-reduce" % world
                                                if world > 1 else "1 GPU",
                                                path=names[used_path]),
            "e2e": {"value": evals_total / (e2e_ms * 1e-3), "unit": UNIT, "ms_per_step": e2e_ms,
                    "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "api": "pmb_batch_prepare/execute/fetch via ctypes (pymasc_b200._shim)",
                    "h2d": ("pipelined: up to %d chromosome groups (none under 16 MiB) on a copy stream"
                            % args.h2d_groups)
                           if args.h2d_groups > 0 else "copy, then compute",
                    "pcie_floor_ms": h2d / 55e9 * 1e3,
                    "note": "PCIe-bound: the step's inputs cross the bus once per step"},
            "gpu_launches": int(launches_per_step * args.steps),
            "launches_per_step": int(launches_per_step),
            "clocks": clocks,
            "roofline": roof,
            "paths": paths,
            "execute_only_ms_per_step_over_ranks": exec_only_ms,
            "resident_equals_e2e": bool(same),
            "synthetic_generation_s": round(gen_s, 1),
        }
    # ---- the other BASELINE configs at this N, and the "s per BAM" half of the metric
    blocks = {}
    if not args.no_configs:
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import bench_blocks
        comm = {"reduce": reduce, "barrier": barrier, "max_over_ranks": max_over_ranks,
                "all_true": lambda ok: max_over_ranks(0.0 if ok else 1.0) == 0.0}
        for key, fn, est_s in (("configs[3]", bench_blocks.deep_library_block, 60.0),
                               ("configs[4]", bench_blocks.multi_file_block, 60.0)):
            spent = max_over_ranks(time.time() - T_START)       # one decision for all ranks
            if spent + est_s > args.budget_s:
                blocks[key] = {"skipped": "time budget: %.0f s spent of %.0f" % (spent, args.budget_s)}
                continue
            with torch.cuda.stream(stream):
                blocks[key] = fn(dev, stream, world, rank, comm, steps=max(3, min(args.steps, 5)))
    if rank == 0:
        line["configs"] = blocks
        # one flag for every full-size consistency check of this run (a False here voids the numbers)
        checks = [bool(same)] + [bool(o["same"]) for o in others.values()] + \
                 [bool(v["resident_equals_e2e"]) and bool(v["counters_equal_oracle_on_sample"])
                  for v in blocks.values() if "ms_per_step" in v]
        line["parity_ok"] = all(checks)
        if not line["parity_ok"]:
            sys.stderr.write("bench.py: PARITY CHECK FAILED (resident / e2e / other paths / oracle sample disagree)\n")
        if world == 1 and not args.no_cpu_baseline:
            dev.set_stream(None)
            line["cpu_baseline"] = cpu_baseline_leg(workload, mode, check_device=dev)
        else:
            line["cpu_baseline"] = None
        if world == 1 and not args.no_bam and time.time() - T_START + 60.0 <= args.budget_s + 120.0:
            sys.path.insert(0, os.path.join(ROOT, "tools"))
            import bench_blocks
            dev.set_stream(None)
            bam = bench_blocks.s_per_bam_block(workload, MAX_SHIFT)
            cb = line.get("cpu_baseline") or {}
            if cb.get("value"):
                cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
                bam["reference_correlation_stage_s"] = {
                    "one_core": evals_total / cb["value"], "all_%d_cores_if_p_N_scaled_perfectly" % cores:
                    evals_total / cb["value"] / cores,
                    "source": "cpu_baseline of this line (all of chr21 on one core), scaled by evals; "
                              "`bench.py --impl reference` measures the -p N rate on all cores"}
            line["s_per_bam"] = bam
        sys.stdout.flush()
        os.dup2(real_stdout, 1)
        print(json.dumps(line), flush=True)
        os.dup2(2, 1)
    dev.set_stream(None)
    dev.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--mode", default="both", choices=["both", "ncc"])
    ap.add_argument("--variant", type=int, default=None, help="shift-window kernel variant (tuning)")
    ap.add_argument("--tile-words", type=int, default=0)
    ap.add_argument("--single-path", action="store_true", help="skip the cross-check pass of the other algorithm")
    ap.add_argument("--path", default="auto", choices=["auto", "dense", "events", "lists"],
                    help="device algorithm: dense shift-window kernel or event-window kernels")
    ap.add_argument("--h2d-groups", type=int, default=4,
                    help="e2e arm: chromosome groups whose H2D overlaps the kernels (0 = no overlap)")
    ap.add_argument("--opt", action="append", default=[], help="pmb_set_option name=value (tuning)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-clocks", action="store_true", help="do not poll NVML during the timed region (diagnostic)")
    ap.add_argument("--no-configs", action="store_true", help="skip the configs[3] / configs[4] blocks")
    ap.add_argument("--no-bam", action="store_true", help="skip the s-per-BAM block (N = 1 only)")
    ap.add_argument("--budget-s", type=float, default=200.0,
                    help="wall-clock budget: a configs block that would start later than this is skipped")
    ap.add_argument("--ref-cores", type=int, default=0)
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    if args.impl == "reference":
        if env_int("RANK", 0) == 0:
            reference_arm(args)
        return
    if args.warmup < 3:
        args.warmup = 3      # timing rule: at least 3 warm-up steps
    gpu_arm(args)


if __name__ == "__main__":
    main()
