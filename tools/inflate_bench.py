#!/usr/bin/env python
"""Single-thread throughput of the BGZF block decoders in libpymasc_io.so against zlib, on two
kinds of 64 KiB blocks:

  literal-heavy   what real BAM payloads look like: records with random read names, packed bases
                  (random bytes) and skewed quality strings, zlib level 6 (compression ratio ~2)
  match-heavy     the synthetic bench file of tools/bam_e2e.py: constant names / bases /
                  qualities, zlib level 1 (ratio ~18)

    python tools/inflate_bench.py [--blocks 48] [--reps 30]

Times pmio_inflate_raw (own decoder, one stream), pmio_inflate_raw2 (own decoder, two streams
with interleaved loops), zlib through the same entry point, and the CRC-32 routes.  Best of
--reps passes over the corpus (the build container's clock is noisy).  Needs no GPU.  One JSON line.
"""
import argparse
import ctypes
import json
import os
import struct
import sys
import time
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _raw(data, level):
    co = zlib.compressobj(level, zlib.DEFLATED, -15)
    return co.compress(data) + co.flush()


def literal_heavy_blocks(n, rng):
    quals = np.array([30, 31, 32, 33, 34, 35, 36, 37, 38, 39, 40, 2], dtype=np.uint8)
    weights = [.02, .03, .05, .05, .08, .1, .12, .15, .15, .12, .1, .03]
    pos, blocks = 1000, []
    for _ in range(n):
        parts = []
        for _ in range(600):
            pos += int(rng.integers(0, 50))
            name = b"SRR1234567.%d" % rng.integers(1, 10 ** 8)
            core = struct.pack("<iiBBHHHiiii", 0, pos, len(name) + 1, 30, 4680, 1, 16 if rng.random() < .5 else 0,
                               36, -1, -1, 0)
            body = (core + name + b"\0" + struct.pack("<I", 36 << 4) + rng.integers(0, 256, 18, dtype=np.uint8).tobytes()
                    + rng.choice(quals, 36, p=weights).tobytes())
            parts.append(struct.pack("<i", len(body)) + body)
        blocks.append(b"".join(parts)[:65280])
    return blocks, 6


def match_heavy_blocks(n, rng):
    rec = np.dtype([("block_size", "<i4"), ("ref", "<i4"), ("pos", "<i4"), ("l_name", "u1"), ("mapq", "u1"),
                    ("bin", "<u2"), ("n_cigar", "<u2"), ("flag", "<u2"), ("l_seq", "<i4"), ("nref", "<i4"),
                    ("npos", "<i4"), ("tlen", "<i4"), ("name", "S2"), ("cigar", "<u4"), ("seq", "S18"), ("qual", "S36")])
    per = 65280 // rec.itemsize
    blocks, pos = [], 1000
    for _ in range(n):
        r = np.zeros(per, rec)
        p = pos + np.cumsum(rng.integers(0, 300, per))
        pos = int(p[-1])
        r["block_size"], r["pos"], r["l_name"], r["mapq"], r["bin"], r["n_cigar"] = rec.itemsize - 4, p, 2, 30, 4680, 1
        r["flag"] = np.where(rng.random(per) < .5, 16, 0)
        r["l_seq"], r["nref"], r["npos"], r["name"], r["cigar"] = 36, -1, -1, b"r", 36 << 4
        r["seq"], r["qual"] = b"\x11" * 18, b"\x28" * 36
        blocks.append(r.tobytes())
    return blocks, 1


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--blocks", type=int, default=48)
    ap.add_argument("--reps", type=int, default=30)
    args = ap.parse_args()
    from pymasc_b200 import io_native
    L = io_native.load_library()
    one = L.pmio_inflate_raw
    one.argtypes = [ctypes.c_char_p, ctypes.c_uint64, ctypes.c_void_p, ctypes.c_uint64, ctypes.c_int]
    two = L.pmio_inflate_raw2
    out_a, out_b = ctypes.create_string_buffer(65536), ctypes.create_string_buffer(65536)
    out_ptr = (ctypes.c_void_p * 2)(ctypes.addressof(out_a), ctypes.addressof(out_b))
    rng = np.random.default_rng(5)
    result = {"metric": "raw-DEFLATE decode of 64 KiB BGZF payloads, one thread", "unit": "MB/s of inflated data",
              "blocks": args.blocks, "reps": args.reps, "cores": len(os.sched_getaffinity(0))}
    for kind, make in (("literal_heavy", literal_heavy_blocks), ("match_heavy", match_heavy_blocks)):
        blocks, level = make(args.blocks, rng)
        comp = [_raw(b, level) for b in blocks]
        total = sum(map(len, blocks))
        pairs = []
        for i in range(0, len(blocks) - 1, 2):
            pairs.append(((ctypes.c_char_p * 2)(comp[i], comp[i + 1]), (ctypes.c_uint64 * 2)(len(comp[i]), len(comp[i + 1])),
                          (ctypes.c_uint64 * 2)(len(blocks[i]), len(blocks[i + 1])), (ctypes.c_int * 2)(0, 0), i))
        best = {"own": 1e9, "own_two_streams": 1e9, "zlib": 1e9, "crc32_clmul": 1e9, "crc32_zlib": 1e9}
        for _ in range(args.reps):
            for label, z in (("own", 0), ("zlib", 1)):
                t = time.perf_counter()
                for c, b in zip(comp, blocks):
                    if one(c, len(c), out_a, len(b), z) != 0:
                        raise SystemExit("decode failed")
                best[label] = min(best[label], time.perf_counter() - t)
            t = time.perf_counter()
            for ins, il, ol, ok, i in pairs:
                two(ins, il, out_ptr, ol, ok)
                if not (ok[0] and ok[1]):
                    raise SystemExit("two-stream decode failed")
            best["own_two_streams"] = min(best["own_two_streams"], time.perf_counter() - t)
            for label, z in (("crc32_clmul", 0), ("crc32_zlib", 1)):
                t = time.perf_counter()
                for b in blocks:
                    L.pmio_crc32(b, len(b), z)
                best[label] = min(best[label], time.perf_counter() - t)
        ins, il, ol, ok, i = pairs[-1]
        two(ins, il, out_ptr, ol, ok)
        assert out_a.raw[:len(blocks[i])] == blocks[i] and out_b.raw[:len(blocks[i + 1])] == blocks[i + 1]
        paired_total = sum(len(blocks[p[4]]) + len(blocks[p[4] + 1]) for p in pairs)
        result[kind] = {"compression_ratio": round(total / sum(map(len, comp)), 2), "zlib_level": level,
                        "own": round(total / best["own"] / 1e6), "own_two_streams": round(paired_total / best["own_two_streams"] / 1e6),
                        "zlib": round(total / best["zlib"] / 1e6), "crc32_clmul": round(total / best["crc32_clmul"] / 1e6),
                        "crc32_zlib": round(total / best["crc32_zlib"] / 1e6)}
    print(json.dumps(result), flush=True)


if __name__ == "__main__":
    main()
