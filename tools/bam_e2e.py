#!/usr/bin/env python
""""s per BAM" (BASELINE.json metric, second half): a synthetic hg38-shaped BAM file and BigWig
written to disk, then timed from the FILES to the genome-wide result objects through
pymasc_b200.pipeline.run_files (native BGZF/BigWig decode -> feed_batch -> B200).

    python tools/bam_e2e.py [--reads 20000000] [--max-shift 1000] [--files 1] [--dir /tmp/pmb]

The writers here are bench infrastructure (vectorised numpy record packing + zlib on a thread
pool); only run_files is timed.  Prints one JSON line.
"""
import argparse
import json
import os
import struct
import sys
import time
import zlib
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

READ_LEN = 36
REC = np.dtype([("block_size", "<i4"), ("ref", "<i4"), ("pos", "<i4"), ("l_name", "u1"), ("mapq", "u1"),
                ("bin", "<u2"), ("n_cigar", "<u2"), ("flag", "<u2"), ("l_seq", "<i4"), ("nref", "<i4"),
                ("npos", "<i4"), ("tlen", "<i4"), ("name", "S2"), ("cigar", "<u4"),
                ("seq", "S%d" % ((READ_LEN + 1) // 2)), ("qual", "S%d" % READ_LEN)])


def _bgzf(payload):
    co = zlib.compressobj(1, zlib.DEFLATED, -15)
    data = co.compress(payload) + co.flush()
    head = struct.pack("<BBBBIBBH", 31, 139, 8, 4, 0, 0, 255, 6) + b"BC" + struct.pack("<HH", 2, 25 + len(data))
    return head + data + struct.pack("<II", zlib.crc32(payload) & 0xFFFFFFFF, len(payload))


def write_bam(path, workload, threads):
    """One BAM record per read of the workload (single-end, MAPQ 30, CIGAR 36M), coordinate
    sorted; returns the number of records."""
    text = b"@HD\tVN:1.6\tSO:coordinate\n"
    head = b"BAM\1" + struct.pack("<i", len(text)) + text + struct.pack("<i", len(workload))
    for w in workload:
        nm = w["name"].encode() + b"\0"
        head += struct.pack("<i", len(nm)) + nm + struct.pack("<i", w["length"])
    n_total = 0
    # like htslib: the header is flushed in blocks of its own and no record straddles two blocks
    # (bgzf_flush_try before a record that would not fit)
    BLK = 65280 // REC.itemsize * REC.itemsize
    with open(path, "wb") as fh, ThreadPoolExecutor(threads) as pool:
        for o in range(0, len(head), 65280):
            fh.write(_bgzf(head[o:o + 65280]))
        pending = b""
        for ci, w in enumerate(workload):
            left = np.concatenate([w["fwd"].astype(np.int64), w["rev"].astype(np.int64) - READ_LEN + 1])
            rev = np.concatenate([np.zeros(w["fwd"].size, bool), np.ones(w["rev"].size, bool)])
            order = np.argsort(left, kind="stable")
            left, rev = left[order], rev[order]
            r = np.zeros(left.size, REC)
            r["block_size"] = REC.itemsize - 4
            r["ref"] = ci
            r["pos"] = left - 1
            r["l_name"], r["mapq"], r["bin"], r["n_cigar"] = 2, 30, 4680, 1
            r["flag"] = np.where(rev, 16, 0)
            r["l_seq"], r["nref"], r["npos"] = READ_LEN, -1, -1
            r["name"] = b"r"
            r["cigar"] = READ_LEN << 4
            r["seq"] = b"\x11" * ((READ_LEN + 1) // 2)
            r["qual"] = b"\x28" * READ_LEN
            n_total += left.size
            stream = pending + r.tobytes()
            cut = len(stream) - len(stream) % BLK
            for blk in pool.map(_bgzf, (stream[o:o + BLK] for o in range(0, cut, BLK))):
                fh.write(blk)
            pending = stream[cut:]
        if pending:
            fh.write(_bgzf(pending))
        fh.write(_bgzf(b""))
    return n_total


def write_bigwig(path, workload):
    import io_writers
    sections = []
    for ci, w in enumerate(workload):
        b, e = w["begins"].tolist(), w["ends"].tolist()
        for o in range(0, len(b), 512):
            sections.append({"chrom_id": ci, "type": "bedgraph",
                             "items": [(x, y, 1.0) for x, y in zip(b[o:o + 512], e[o:o + 512])]})
    io_writers.write_bigwig(path, [(w["name"], w["length"]) for w in workload], sections)
    return sum(len(w["begins"]) for w in workload)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reads", type=int, default=20_000_000)
    ap.add_argument("--max-shift", type=int, default=1000)
    ap.add_argument("--files", type=int, default=1)
    ap.add_argument("--dir", default="/tmp/pmb_e2e")
    ap.add_argument("--threads", type=int, default=0)
    ap.add_argument("--repeat", type=int, default=2)
    ap.add_argument("--chroms", default="", help="comma-separated subset (tests)")
    ap.add_argument("--oracle-engine", action="store_true", help="CPU check of the tool itself (tiny inputs)")
    args = ap.parse_args()
    from pymasc_b200 import pipeline, synth
    os.makedirs(args.dir, exist_ok=True)
    cores = len(os.sched_getaffinity(0))
    threads = args.threads or cores
    chroms = [c for c in args.chroms.split(",") if c] or None
    t0 = time.time()
    bams = []
    n_reads = 0
    wl = None
    for f in range(args.files):
        wl = synth.workload(n_reads=args.reads, read_len=READ_LEN, chroms=chroms, seed=20260101 + 100 * f)
        p = os.path.join(args.dir, "synth_%d.bam" % f)
        n_reads += write_bam(p, wl, threads)
        bams.append(p)
    bw = os.path.join(args.dir, "synth_36mer.bigwig")
    n_iv = write_bigwig(bw, wl)
    gen_s = time.time() - t0
    engine = None
    if args.oracle_engine:
        from oracle_engine import OracleEngine
        engine = OracleEngine()
    runs = []
    for _ in range(args.repeat):
        t0 = time.perf_counter()
        results, info = pipeline.run_files(bams, bw, max_shift=args.max_shift, mapq_criteria=1,
                                           threads=threads, engine=engine)
        info["seconds"]["wall"] = time.perf_counter() - t0
        runs.append(info["seconds"])
    best = min(runs, key=lambda s: s["wall"])
    whole = results[0]
    evals = synth.evals([w["length"] for w in wl], READ_LEN, args.max_shift) * args.files
    line = {
        "metric": "s per BAM (files on disk -> genome-wide NCC+MSCC result objects)", "unit": "s",
        "value": best["wall"] / args.files, "higher_is_better": False, "files": args.files,
        "reads_in_files": int(n_reads), "intervals": int(n_iv), "max_shift": args.max_shift,
        "bam_bytes": sum(os.path.getsize(p) for p in bams), "bigwig_bytes": os.path.getsize(bw),
        "host_threads": threads, "stages_s": best, "all_runs": runs,
        "evals": evals, "evals_per_s_from_files": evals / best["wall"],
        "read_len_estimated": info["files"][0]["read_len"],
        "result": {"type": type(whole).__name__, "forward_sum": int(whole.forward_sum),
                   "reverse_sum": int(whole.reverse_sum), "chroms": len(whole.chroms)},
        "file_generation_s": round(gen_s, 1),
    }
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
