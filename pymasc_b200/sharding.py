"""Multi-GPU partitioning of the cross-correlation counts (host logic, no device code).

The reference parallelises by chromosome only (one OS process per chromosome,
PyMaSC/handler/calc.py:163-192, PyMaSC/utils/calc.py:119-130).  Every count of
``_calc_correlation`` (PyMaSC/core/bitarray/mscc.pyx:288-317) is a sum over bit positions j,
so a chromosome can also be cut into position ranges [lo, hi) whose partial count vectors
add up exactly:

    rank r counts j in [lo_r, hi_r); it needs F on [lo, hi), R on [lo, hi + max_shift] and
    M on [lo - max_shift, hi + read_len) -- the halos are rebuilt on each GPU from the same
    packed position / interval arrays (no device-to-device halo exchange), and strand totals
    are counted by bit index so a halo bit is never counted twice.

``plan`` cuts the concatenation of all chromosomes' bit arrays into `world` equal contiguous
pieces (so ranks are balanced to within one alignment unit whatever the chromosome sizes are);
``jobs_for_rank`` turns a rank's pieces into C-ABI jobs over an identical job list on every
rank (chromosomes outside the rank get an empty range), so every rank's device counter block
has the same layout and ONE all-reduce (int64 sum, NCCL over NVLink) completes the result.
"""
import numpy as np

from . import _shim

ALIGN_BITS = 128   # piece boundaries are multiples of 128 bits (one 16-byte TMA granule)


def _first_at_least(a, x):
    """Index of the first element of the sorted array `a` that is >= x.  The key is cast to the
    array's own dtype: with a Python int numpy would convert the whole array to int64 on every
    call (0.24 ms per lookup on a 1 M-element uint32 array instead of a microsecond)."""
    if a.dtype.kind in "ui":
        info = np.iinfo(a.dtype)
        if x > info.max:
            return len(a)
        x = a.dtype.type(max(x, info.min))
    return int(np.searchsorted(a, x, side="left"))


def array_bits(chrom_len, read_len, max_shift):
    """Length of a chromosome's bit arrays (mscc.pyx:134, :165)."""
    return int(chrom_len) + int(read_len) + int(max_shift) + 100


def plan(lengths, read_len, max_shift, world):
    """[[(chrom_index, lo, hi), ...] for each rank]: contiguous equal cuts of the genome.

    Every bit j of every chromosome array belongs to exactly one rank."""
    nbits = [array_bits(L, read_len, max_shift) for L in lengths]
    total = sum(nbits)
    cuts = [0]
    for r in range(1, world):
        c = (total * r // world) // ALIGN_BITS * ALIGN_BITS
        cuts.append(max(c, cuts[-1]))
    cuts.append(total)
    return _pieces_from_cuts(cuts, nbits, world)


def _pieces_from_cuts(cuts, nbits, world):
    out = [[] for _ in range(world)]
    base = 0
    for ci, nb in enumerate(nbits):
        for r in range(world):
            lo = max(cuts[r], base) - base
            hi = min(cuts[r + 1], base + nb) - base
            if hi > lo:
                out[r].append((ci, lo, hi))
        base += nb
    return out


# Relative cost of one read / one run edge of the mappability track against one bit position, for
# the event-window kernels (ncu instruction counts, profiles/r01_ncu_event_window_v9.txt: ~0.4
# warp instructions per bit position over the build and the three scans, ~60 per event).
READ_COST = 150.0
EDGE_COST = 150.0


def plan_weighted(chroms, read_len, max_shift, world, read_cost=READ_COST, edge_cost=EDGE_COST):
    """Like ``plan``, but the cuts equalise bits + read_cost * reads + edge_cost * run edges per
    rank (a rank's time follows its events as much as its length).  `chroms`: dicts with
    ``length``, ``fwd``, ``rev`` and optionally ``begins`` (sorted arrays, as for
    ``jobs_for_rank``); every rank must call it with the same data."""
    nbits = [array_bits(c["length"], read_len, max_shift) for c in chroms]

    def arrays(c):
        out = [(a, read_cost) for a in (c.get("fwd"), c.get("rev")) if a is not None and len(a)]
        b = c.get("begins")
        if b is not None and len(b):
            out.append((b, 2.0 * edge_cost))
        return out

    def cost_upto(c, x):        # cost of bits [0, x) of chromosome c
        return float(x) + sum(w * float(_first_at_least(a, x)) for a, w in arrays(c))

    whole = [cost_upto(c, nb) for c, nb in zip(chroms, nbits)]
    total = sum(whole)
    bases = np.concatenate([[0], np.cumsum(nbits)]).astype(np.int64)
    cuts = [0]
    for r in range(1, world):
        target = total * r / world
        acc, ci = 0.0, 0
        while ci < len(chroms) - 1 and acc + whole[ci] < target:
            acc += whole[ci]
            ci += 1
        lo, hi = 0, nbits[ci]                      # first x with cost_upto >= target - acc
        while lo < hi:
            mid = (lo + hi) // 2
            if acc + cost_upto(chroms[ci], mid) < target:
                lo = mid + 1
            else:
                hi = mid
        cut = (int(bases[ci]) + lo) // ALIGN_BITS * ALIGN_BITS
        cuts.append(max(cut, cuts[-1]))
    cuts.append(int(bases[-1]))
    return _pieces_from_cuts(cuts, nbits, world)


def plan_by_chromosome(lengths, read_len, max_shift, world):
    """Whole chromosomes only, longest-processing-time-first bin packing (the reference's
    -p N unit of work).  Same return shape as ``plan``."""
    nbits = [array_bits(L, read_len, max_shift) for L in lengths]
    load = [0] * world
    out = [[] for _ in range(world)]
    for ci in sorted(range(len(nbits)), key=lambda i: -nbits[i]):
        r = min(range(world), key=lambda i: load[i])
        load[r] += nbits[ci]
        out[r].append((ci, 0, nbits[ci]))
    for pieces in out:
        pieces.sort()
    return out


def _clip_sorted(a, lo, hi):
    """View of the sorted array `a` restricted to lo <= a < hi."""
    if a is None or len(a) == 0:
        return a
    i0 = _first_at_least(a, max(lo, 0))
    i1 = _first_at_least(a, max(hi, 0))
    return a[i0:i1]


def jobs_for_rank(chroms, pieces, read_len, max_shift, flags_of=None):
    """C-ABI jobs of one rank.

    chroms: list of dicts with keys length, fwd, rev, begins, ends (sorted packed arrays;
    begins/ends None without mappability).  pieces: that rank's entry of ``plan``.
    Returns one job per chromosome, in chromosome order, on EVERY rank; a chromosome the rank
    does not own gets an empty range and no inputs.  A chromosome cut into two pieces on the
    same rank cannot happen (pieces are contiguous), so one job per chromosome suffices.
    """
    by_chrom = {ci: (lo, hi) for ci, lo, hi in pieces}
    jobs = []
    for ci, c in enumerate(chroms):
        has_map = c.get("begins") is not None
        flags = flags_of(c) if flags_of else (_shim.FLAG_NCC | (_shim.FLAG_MSCC if has_map else 0))
        nb = array_bits(c["length"], read_len, max_shift)
        if ci not in by_chrom:
            jobs.append(_shim.Job(c["length"], read_len, max_shift, flags, None, None,
                                  None, None, nb, nb))
            continue
        lo, hi = by_chrom[ci]
        fwd = _clip_sorted(c["fwd"], lo, hi)
        rev = _clip_sorted(c["rev"], lo, hi + max_shift + 1)
        b = e = None
        if has_map and flags & (_shim.FLAG_MSCC | _shim.FLAG_AUTOCORR):
            # interval (b, e) sets bits b+1..e; keep those touching the shard's M window:
            # [lo - max_shift, hi + read_len) for D_k, [lo, hi + max_shift] for the autocorrelation
            ends, begins = c["ends"], c["begins"]
            i0 = _first_at_least(ends, max(lo - max_shift - 1, 0))
            i1 = _first_at_least(begins, hi + max(read_len, max_shift + 1))
            b, e = begins[i0:i1], ends[i0:i1]
        whole = lo == 0 and hi == nb
        jobs.append(_shim.Job(c["length"], read_len, max_shift, flags, fwd, rev, b, e,
                              0 if whole else lo, 0 if whole else hi))
    return jobs


def sum_outputs(per_rank_outputs):
    """Host-side equivalent of the all-reduce (tests, single-process multi-device runs):
    element-wise sum of the per-rank output dicts of the same job list."""
    first = per_rank_outputs[0]
    merged = []
    for j, job_out in enumerate(first):
        d = {}
        for k, v0 in job_out.items():
            if np.isscalar(v0) or isinstance(v0, int):
                d[k] = sum(int(outs[j][k]) for outs in per_rank_outputs)
            else:
                acc = np.array(v0, dtype=np.int64, copy=True)
                for outs in per_rank_outputs[1:]:
                    acc += outs[j][k]
                d[k] = acc
        merged.append(d)
    return merged
