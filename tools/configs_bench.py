#!/usr/bin/env python
"""BASELINE.json configs[2], [3] and [4] (bench.py covers configs[1]/[2] over position shards).

    python tools/configs_bench.py --config 4 [--reads 200000000] [--steps 5]       # 1 GPU
    python tools/configs_bench.py --config 5 [--files 16] [--steps 5]              # 1 GPU, shared vs per-file track
    python tools/configs_bench.py --config 3 --devices 4     # one genome, whole chromosomes per GPU, 1/2/4 GPUs
    python tools/configs_bench.py --config 5 --devices 4     # 16 read sets, shared-track runs per GPU
    python tools/configs_bench.py --config 4 --devices 4     # deep library, position ranges per GPU
    (--devices runs one process over 1, 2, 4 ... GPUs through pymasc_b200.multigpu.MultiDeviceEngine;
     PMB_BREAKDOWN=1 prints the host-phase breakdown of the 1-GPU end-to-end step)

config 4  deep library: 200 M reads, NCC+MSCC, max_shift 5000 (evals = 1.544e13).
config 5  16 read sets (seeds 20260101 + 100 b + chrom_index) sharing the config-3 mappability
          track, MSCC only (skip_ncc), max_shift 1000.  One device batch of 16 x 24 jobs in
          chromosome-major order; jobs 2..16 of a chromosome carry PMB_FLAG_SAME_TRACK, so M,
          Er/Or and dmp are built once per chromosome.  The same batch with every job carrying
          its own copy of the intervals is timed beside it and must give identical counters.

Prints one JSON line per config: device time per step with resident inputs (CUDA events through
the library's own timing + torch events on the launching stream), and end to end from pinned
host arrays (H2D + build + kernels + D2H).
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

READ_LEN = 36


def timed(dev, stream, jobs, steps, warmup, e2e_groups=4):
    """(resident ms/step, e2e ms/step, outputs, timing dict, h2d bytes) of one job list."""
    import torch
    import bench
    h2d = bench.pinned_pack(jobs)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        batch = dev.prepare(jobs)
        for _ in range(warmup):
            batch.execute()
        torch.cuda.synchronize()
        ev0.record(stream)
        for _ in range(steps):
            batch.execute()
        ev1.record(stream)
        torch.cuda.synchronize()
        resident_ms = ev0.elapsed_time(ev1) / steps
        batch.execute()
        t = dev.timing()
        outs = batch.fetch()
        batch.free()

        dev.set_option("async_h2d", 1)
        dev.set_option("h2d_groups", e2e_groups)

        def e2e_step():
            b = dev.prepare(jobs)
            try:
                b.execute()
                return b.fetch()
            finally:
                b.free()

        for _ in range(min(warmup, 2)):
            e2e_step()
        torch.cuda.synchronize()
        if os.environ.get("PMB_BREAKDOWN"):     # host wall time per phase of one e2e step (stderr)
            acc = np.zeros(5)
            for _ in range(steps):
                tt = [time.perf_counter()]
                b = dev.prepare(jobs); tt.append(time.perf_counter())
                b.execute(); tt.append(time.perf_counter())
                dev.synchronize(); tt.append(time.perf_counter())
                b.fetch(); tt.append(time.perf_counter())
                b.free(); tt.append(time.perf_counter())
                acc += np.diff(tt)
            print("[e2e breakdown] prepare %.2f execute_call %.2f wait %.2f fetch %.2f free %.2f ms; "
                  "device execute %.2f ms" % (tuple(acc / steps * 1e3) + (dev.timing()["total_ms"],)),
                  file=sys.stderr)
        t0 = time.perf_counter()
        for _ in range(steps):
            e2e_out = e2e_step()
        torch.cuda.synchronize()
        e2e_ms = (time.perf_counter() - t0) * 1e3 / steps
        dev.set_option("async_h2d", 0)
    same = all(np.array_equal(np.asarray(a[k]), np.asarray(b[k])) for a, b in zip(outs, e2e_out) for k in a)
    d2h = sum(sum(v.size * 8 for v in o.values() if isinstance(v, np.ndarray)) + 16 for o in outs)
    return resident_ms, e2e_ms, outs, t, h2d, d2h, same


def config4(args, dev, stream):
    from pymasc_b200 import _shim, synth
    D = args.max_shift or 5000
    t0 = time.time()
    wl = synth.workload(n_reads=args.reads, read_len=READ_LEN)
    gen_s = time.time() - t0
    jobs = [_shim.Job(w["length"], READ_LEN, D, 3, w["fwd"], w["rev"], w["begins"], w["ends"]) for w in wl]
    evals = synth.evals([w["length"] for w in wl], READ_LEN, D)
    res_ms, e2e_ms, outs, t, h2d, d2h, same = timed(dev, stream, jobs, args.steps, args.warmup)
    # size-independent checks at full size: popc totals, k = 0 identity, dmp symmetry
    ok = True
    for w, o in zip(wl, outs):
        ok &= o["fwd_total"] == w["fwd"].size and o["rev_total"] == w["rev"].size
        ok &= int(o["ncc"][0]) == np.intersect1d(w["fwd"], w["rev"]).size
        ok &= all(o["dmp"][k] == o["dmp"][2 * (READ_LEN - 1) - k] for k in range(READ_LEN))
        ok &= bool(np.all(o["mscc"] <= np.minimum(o["mfwd"], o["mrev"])))
    n_reads = sum(w["fwd"].size + w["rev"].size for w in wl)
    return {
        "config": "BASELINE configs[3]: deep library, NCC+MSCC, 1 B200 (whole genome, no sharding)",
        "metric": "bit-position x shift evals/s", "unit": "evals/s", "n_gpus": 1,
        "reads_requested": args.reads, "distinct_5prime_positions": int(n_reads), "max_shift": D,
        "evals_per_step": evals, "ms_per_step": res_ms, "value": evals / (res_ms * 1e-3),
        "e2e": {"ms_per_step": e2e_ms, "value": evals / (e2e_ms * 1e-3), "h2d_bytes_per_step": int(h2d),
                "d2h_bytes_per_step": int(d2h)},
        "timing": t, "properties_hold": bool(ok), "resident_equals_e2e": bool(same),
        "steps": args.steps, "warmup": args.warmup, "synthetic_generation_s": round(gen_s, 1),
    }


def config5(args, dev, stream):
    from pymasc_b200 import _shim, synth
    D = args.max_shift or 1000
    S = _shim.FLAG_SAME_TRACK
    t0 = time.time()
    names = [n for n, _ in synth.HG38]
    lengths = [L for _, L in synth.HG38]
    total = float(sum(lengths))
    track = [synth.mappability(i, L) for i, L in enumerate(lengths)]
    shared, private = [], []
    for i, L in enumerate(lengths):                 # chromosome-major: the files of one chromosome are a run
        n = max(8, int(round(20_000_000 * L / total)))
        for f in range(args.files):
            fwd, rev = synth.chip_reads(i, L, n, READ_LEN, seed=20260101 + 100 * f)
            b, e = track[i]
            shared.append(_shim.Job(L, READ_LEN, D, _shim.FLAG_MSCC | (S if f else 0), fwd, rev,
                                    None if f else b, None if f else e))
            private.append(_shim.Job(L, READ_LEN, D, _shim.FLAG_MSCC, fwd, rev, b.copy(), e.copy()))
    gen_s = time.time() - t0
    evals = synth.evals(lengths, READ_LEN, D) * args.files
    out = {
        "config": "BASELINE configs[4]: %d read sets sharing one mappability track, MSCC only, "
                  "max_shift %d, 1 B200 (one batch of %d jobs)" % (args.files, D, len(shared)),
        "metric": "bit-position x shift evals/s", "unit": "evals/s", "n_gpus": 1, "files": args.files,
        "evals_per_step": evals, "steps": args.steps, "warmup": args.warmup,
        "synthetic_generation_s": round(gen_s, 1),
    }
    res = {}
    for name, jobs in (("shared_track", shared), ("track_per_file", private)):
        res_ms, e2e_ms, outs, t, h2d, d2h, same = timed(dev, stream, jobs, args.steps, args.warmup)
        res[name] = outs
        out[name] = {"ms_per_step": res_ms, "ms_per_file": res_ms / args.files, "value": evals / (res_ms * 1e-3),
                     "e2e": {"ms_per_step": e2e_ms, "value": evals / (e2e_ms * 1e-3),
                             "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h)},
                     "timing": t, "resident_equals_e2e": bool(same)}
    out["counters_equal"] = all(np.array_equal(np.asarray(a[k]), np.asarray(b[k]))
                                for a, b in zip(res["shared_track"], res["track_per_file"]) for k in a)
    out["ms_per_step"] = out["shared_track"]["ms_per_step"]
    out["value"] = out["shared_track"]["value"]
    return out


def by_chromosome(args):
    """configs[2] (one genome) or configs[4] (--files read sets over one track) on --devices GPUs
    of the box, one process: whole chromosomes per GPU (the reference's -p N unit), no
    collective.  Timed end to end from pinned host arrays, wall clock around
    MultiDeviceEngine.run, with each device's own CUDA-event time beside it."""
    import torch
    import bench
    from pymasc_b200 import _shim, synth
    from pymasc_b200.multigpu import MultiDeviceEngine
    D = args.max_shift or 1000
    S = _shim.FLAG_SAME_TRACK
    lengths = [L for _, L in synth.HG38]
    total = float(sum(lengths))
    flags = _shim.FLAG_MSCC | (0 if args.files > 1 else _shim.FLAG_NCC)
    t0 = time.time()
    jobs = []
    for i, L in enumerate(lengths):
        n = max(8, int(round(20_000_000 * L / total)))
        b, e = synth.mappability(i, L)
        for f in range(args.files):
            fwd, rev = synth.chip_reads(i, L, n, READ_LEN, seed=20260101 + 100 * f)
            jobs.append(_shim.Job(L, READ_LEN, D, flags | (S if f else 0), fwd, rev,
                                  None if f else b, None if f else e))
    gen_s = time.time() - t0
    h2d = bench.pinned_pack(jobs)
    evals = synth.evals(lengths, READ_LEN, D) * args.files
    out = {"config": "%d read set(s), %s, max_shift %d, whole chromosomes per GPU, one process" %
                     (args.files, "MSCC only, shared track" if args.files > 1 else "NCC+MSCC", D),
           "metric": "bit-position x shift evals/s", "unit": "evals/s", "evals_per_step": evals,
           "h2d_bytes_per_step": int(h2d), "steps": args.steps, "warmup": args.warmup,
           "synthetic_generation_s": round(gen_s, 1), "by_devices": {}}
    ref = None
    n_dev = 1
    while n_dev <= args.devices:
        eng = MultiDeviceEngine(list(range(n_dev)))
        for _ in range(args.warmup):
            res = eng.run(jobs)
        torch.cuda.synchronize()
        walls, dev_ms = [], []
        for _ in range(args.steps):
            t0 = time.perf_counter()
            res = eng.run(jobs)
            walls.append((time.perf_counter() - t0) * 1e3)
            dev_ms.append(max(x for x in eng.last_device_ms if x is not None))
        eng.close()
        if ref is None:
            ref = res
        same = all(np.array_equal(np.asarray(a[k]), np.asarray(b2[k])) for a, b2 in zip(ref, res) for k in a)
        out["by_devices"][str(n_dev)] = {
            "e2e_ms_per_step": float(np.mean(walls)), "e2e_ms_min": float(np.min(walls)),
            "value_e2e": evals / (float(np.mean(walls)) * 1e-3),
            "slowest_device_execute_ms": float(np.mean(dev_ms)),
            "jobs_per_device": [len(idx) for idx in eng.last_placement],
            "equals_1_device": bool(same)}
        n_dev *= 2
    return out


def by_position(args):
    """configs[3] (deep library, max_shift 5000) on --devices GPUs, one process: every GPU counts
    one position range of the concatenated genome with halos rebuilt from its clipped inputs
    (MultiDeviceEngine(split="position")), count vectors summed on the host.  End to end from
    pinned host arrays, wall clock around run()."""
    import torch
    import bench
    from concurrent.futures import ThreadPoolExecutor
    from pymasc_b200 import _shim, synth
    from pymasc_b200.multigpu import MultiDeviceEngine
    D = args.max_shift or 5000
    lengths = [L for _, L in synth.HG38]
    total = float(sum(lengths))
    t0 = time.time()

    def make(i):
        L = lengths[i]
        fwd, rev = synth.chip_reads(i, L, max(8, int(round(args.reads * L / total))), READ_LEN)
        b, e = synth.mappability(i, L)
        return _shim.Job(L, READ_LEN, D, 3, fwd, rev, b, e)
    with ThreadPoolExecutor(8) as pool:
        jobs = list(pool.map(make, range(len(lengths))))
    gen_s = time.time() - t0
    h2d = bench.pinned_pack(jobs)
    evals = synth.evals(lengths, READ_LEN, D)
    out = {"config": "BASELINE configs[3]: deep library (%d reads requested, %d distinct 5' positions), NCC+MSCC, "
                     "max_shift %d, position ranges with halos, one process" %
                     (args.reads, sum(j.fwd_pos.size + j.rev_pos.size for j in jobs), D),
           "metric": "bit-position x shift evals/s", "unit": "evals/s", "evals_per_step": evals,
           "h2d_bytes_whole_genome": int(h2d), "steps": args.steps, "warmup": args.warmup,
           "synthetic_generation_s": round(gen_s, 1), "by_devices": {}}
    ref = None
    n_dev = 1
    while n_dev <= args.devices:
        eng = MultiDeviceEngine(list(range(n_dev)), split="position")
        for _ in range(args.warmup):
            res = eng.run(jobs)
        torch.cuda.synchronize()
        walls, dev_ms = [], []
        for _ in range(args.steps):
            t0 = time.perf_counter()
            res = eng.run(jobs)
            walls.append((time.perf_counter() - t0) * 1e3)
            dev_ms.append([x for x in eng.last_device_ms])
        eng.close()
        if ref is None:
            ref = res
        same = all(np.array_equal(np.asarray(a[k]), np.asarray(b2[k])) for a, b2 in zip(ref, res) for k in a)
        out["by_devices"][str(n_dev)] = {
            "e2e_ms_per_step": float(np.mean(walls)), "e2e_ms_min": float(np.min(walls)),
            "value_e2e": evals / (float(np.mean(walls)) * 1e-3),
            "device_execute_ms": [float(np.mean([d[k] for d in dev_ms])) for k in range(n_dev)],
            "equals_1_device": bool(same)}
        n_dev *= 2
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", type=int, required=True, choices=[3, 4, 5])
    ap.add_argument("--devices", type=int, default=0,
                    help="config 3 / 5: run on 1, 2, 4 ... up to this many GPUs through MultiDeviceEngine")
    ap.add_argument("--reads", type=int, default=200_000_000)
    ap.add_argument("--files", type=int, default=16)
    ap.add_argument("--max-shift", type=int, default=0)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--path", default="auto", choices=["auto", "dense", "events"])
    args = ap.parse_args()
    import torch
    from pymasc_b200 import _shim
    if args.config == 4 and args.devices:
        print(json.dumps(by_position(args)), flush=True)
        return
    if args.config == 3 or args.devices:
        if args.config == 3:
            args.files = 1
        print(json.dumps(by_chromosome(args)), flush=True)
        return
    torch.cuda.set_device(0)
    dev = _shim.Device(0)
    stream = torch.cuda.Stream(device=0)
    dev.set_stream(stream.cuda_stream)
    dev.set_option("path", {"auto": 0, "dense": 1, "events": 2}[args.path])
    line = (config4 if args.config == 4 else config5)(args, dev, stream)
    dev.set_stream(None)
    dev.close()
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
