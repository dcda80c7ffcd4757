#!/usr/bin/env python
"""Host-only timing of the BAM decoder (libpymasc_io.so): a synthetic hg38-shaped BAM written to
disk, then pymasc_b200.io_native.BamFile timed with the library's own inflate + CRC and with
zlib (PMIO_ZLIB_ONLY=1, a child process each: the switch is read once per process).

    python tools/decode_bench.py [--reads 8000000] [--dir /tmp/pmb_dec] [--threads 0]

Needs no GPU.  Prints one JSON line; the per-stage split (PMIO_TIMING=1) goes to stderr.
"""
import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "tools")):
    if p not in sys.path:
        sys.path.insert(0, p)

CHILD = """
import sys, time, json
sys.path.insert(0, %r)
from pymasc_b200 import io_native
best = None
for _ in range(%d):
    t = time.perf_counter()
    bam = io_native.BamFile(%r, mapq_criteria=1, threads=%d)
    dt = time.perf_counter() - t
    best = dt if best is None else min(best, dt)
    n = bam.stats["n_accepted"]
    bam.close()
print(json.dumps({"s": best, "n_accepted": n}))
"""


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reads", type=int, default=8_000_000)
    ap.add_argument("--dir", default="/tmp/pmb_dec")
    ap.add_argument("--threads", type=int, default=0)
    ap.add_argument("--repeat", type=int, default=3)
    args = ap.parse_args()
    import bam_e2e
    from pymasc_b200 import synth
    os.makedirs(args.dir, exist_ok=True)
    cores = len(os.sched_getaffinity(0))
    threads = args.threads or cores
    t0 = time.time()
    wl = synth.workload(n_reads=args.reads, read_len=bam_e2e.READ_LEN, seed=20260101)
    path = os.path.join(args.dir, "decode.bam")
    n = bam_e2e.write_bam(path, wl, threads)
    gen_s = time.time() - t0
    out = {}
    for label, extra in (("own", {}), ("zlib", {"PMIO_ZLIB_ONLY": "1"})):
        env = {k: v for k, v in os.environ.items() if k != "PMIO_ZLIB_ONLY"}
        env.update(extra, PMIO_TIMING="1")
        r = subprocess.run([sys.executable, "-c", CHILD % (ROOT, args.repeat, path, threads)], env=env,
                           check=True, capture_output=True, text=True)
        sys.stderr.write(r.stderr)
        out[label] = json.loads(r.stdout.strip().splitlines()[-1])
    print(json.dumps({"metric": "BAM decode (file -> packed per-chromosome read arrays)", "unit": "s",
                      "own_inflate_and_crc_s": out["own"]["s"], "zlib_s": out["zlib"]["s"],
                      "records": int(n), "accepted": out["own"]["n_accepted"], "bam_bytes": os.path.getsize(path),
                      "inflated_bytes": int(n) * 76, "host_threads": threads, "cores": cores,
                      "file_generation_s": round(gen_s, 1)}), flush=True)


if __name__ == "__main__":
    main()
