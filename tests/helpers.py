"""Shared test helpers: fixture loading, toy-case generation, result snapshots."""
import json
import math
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def load_fixture():
    """The reference's ChIP-seq fixture as packed arrays (made by tests/golden/make_golden.py
    from tests/data/ENCFF000RMB-test.sam + hg19_36mer-test.bedGraph, MAPQ >= 10)."""
    z = np.load(os.path.join(GOLDEN, "fixture_encff000rmb.npz"), allow_pickle=False)
    return {
        "references": [str(x) for x in z["references"]],
        "lengths": [int(x) for x in z["lengths"]],
        "pos": z["chr1_pos"].astype(np.int64),
        "readlen": z["chr1_readlen"].astype(np.int64),
        "is_reverse": z["chr1_is_reverse"].astype(bool),
        "map_begins": z["chr1_map_begins"].astype(np.uint32),
        "map_ends": z["chr1_map_ends"].astype(np.uint32),
    }


def load_golden_tables():
    with open(os.path.join(GOLDEN, "golden_encff000rmb.json")) as fh:
        return json.load(fh)


def load_ref_vectors():
    with open(os.path.join(GOLDEN, "ref_vectors.json")) as fh:
        return json.load(fh)


class ListFeeder(object):
    """bwfeeder stand-in for the B200 calculator: fetch(thr, chrom) -> iterable of
    (begin, end, value); KeyError when the chromosome has no track
    (PyMaSC/reader/bigwig.pyx:166-167)."""

    def __init__(self, tracks):
        self.tracks = tracks

    def fetch(self, valfilter, chrom):
        if chrom not in self.tracks:
            raise KeyError(chrom)
        b, e = self.tracks[chrom]
        return [(int(x), int(y), 1.0) for x, y in zip(b, e)]


def _clean(x):
    """JSON-friendly, comparison-friendly form of a result field."""
    if x is None:
        return None
    if isinstance(x, (list, tuple, np.ndarray)):
        return [_clean(v) for v in list(x)]
    if isinstance(x, (float, np.floating)):
        return None if math.isnan(float(x)) else float(x)
    if isinstance(x, (int, np.integer)):
        return int(x)
    return x


def snapshot_chrom(res):
    if res is None:
        return None
    d = {"type": type(res).__name__}
    for k in ("genomelen", "max_shift", "read_len", "forward_sum", "reverse_sum",
              "forward_read_len_sum", "reverse_read_len_sum", "ccbins", "mappable_len", "cc"):
        if hasattr(res, k):
            d[k] = _clean(getattr(res, k))
    return d


def snapshot_whole(whole):
    d = {"type": type(whole).__name__}
    for k in ("genomelen", "forward_sum", "reverse_sum", "forward_read_len_sum",
              "reverse_read_len_sum"):
        if hasattr(whole, k):
            d[k] = _clean(getattr(whole, k))
    for k in ("chroms", "mappable_chroms"):
        if hasattr(whole, k):
            d[k] = {c: snapshot_chrom(r) for c, r in getattr(whole, k).items()}
    return d


def assert_snapshots_equal(got, want, path="", rtol=1e-9):
    """Integers and structure exact; floats (cc curves) within rtol (north_star: 1e-9)."""
    if isinstance(want, dict):
        assert isinstance(got, dict), "%s: %r is not a dict" % (path, type(got))
        assert sorted(got) == sorted(want), "%s: keys %s != %s" % (path, sorted(got), sorted(want))
        for k in want:
            assert_snapshots_equal(got[k], want[k], path + "/" + str(k), rtol)
    elif isinstance(want, list):
        assert isinstance(got, list) and len(got) == len(want), \
            "%s: length %s != %s" % (path, len(got) if isinstance(got, list) else got, len(want))
        for i, (g, w) in enumerate(zip(got, want)):
            assert_snapshots_equal(g, w, "%s[%d]" % (path, i), rtol)
    elif isinstance(want, float):
        assert isinstance(got, (int, float)), "%s: %r vs %r" % (path, got, want)
        assert abs(got - want) <= rtol * max(abs(want), 1e-300), "%s: %r != %r" % (path, got, want)
    else:
        assert got == want, "%s: %r != %r" % (path, got, want)


# ----------------------------------------------------------------------------------------
# toy cases (host-semantics coverage: duplicates, variable read lengths, missing tracks,
# read-less chromosomes, skip_ncc, max_shift below/above the read-length thresholds)
# ----------------------------------------------------------------------------------------
TOY_SPECS = [
    # name, seed, chrom lengths, read_len, max_shift, n_reads per chrom, track?, skip_ncc
    dict(name="both_basic", seed=11, lens=[5000, 3000, 4000], rl=20, D=100,
         nreads=[600, 0, 300], tracks=[True, True, False], skip_ncc=False),
    dict(name="mscc_only", seed=12, lens=[5000, 3000, 4000], rl=20, D=100,
         nreads=[600, 0, 300], tracks=[True, True, False], skip_ncc=True),
    dict(name="ncc_only", seed=13, lens=[7001, 2049], rl=36, D=64,
         nreads=[900, 200], tracks=None, skip_ncc=False),
    dict(name="shift_lt_readlen", seed=14, lens=[6000], rl=50, D=30,
         nreads=[800], tracks=[True], skip_ncc=False),
    dict(name="shift_lt_2readlen", seed=15, lens=[6000], rl=50, D=80,
         nreads=[800], tracks=[True], skip_ncc=False),
    dict(name="shift_eq_2readlen_m1", seed=16, lens=[6000, 1500], rl=25, D=49,
         nreads=[700, 100], tracks=[True, True], skip_ncc=False),
    dict(name="one_strand_only", seed=17, lens=[4000], rl=20, D=70,
         nreads=[-400], tracks=[True], skip_ncc=False),   # negative: forward reads only
    dict(name="empty_track", seed=18, lens=[4000, 2500], rl=20, D=70,
         nreads=[400, 100], tracks=["empty", True], skip_ncc=False),
]


def make_toy(spec):
    """Deterministic toy inputs.  Returns dict with references, lengths, reads
    (chrom -> pos/readlen/is_reverse in feed order) and tracks (chrom -> begins, ends)."""
    rng = np.random.default_rng(spec["seed"])
    refs = ["c%d" % (i + 1) for i in range(len(spec["lens"]))]
    reads, tracks = {}, ({} if spec["tracks"] is not None else None)
    for i, (c, L) in enumerate(zip(refs, spec["lens"])):
        n = spec["nreads"][i]
        fwd_only = n < 0
        n = abs(n)
        if n:
            rl = spec["rl"]
            # variable query lengths (some shorter, a few longer than read_len)
            ln = rl + rng.integers(-6, 4, size=n)
            ln = np.clip(ln, 5, None)
            pos = np.sort(rng.integers(1, L - rl - 4, size=n))
            # force exact duplicates and strand collisions
            pos[1::7] = pos[0::7][:len(pos[1::7])]
            pos = np.sort(pos)
            rev = rng.random(n) < 0.5
            if fwd_only:
                rev[:] = False
            # reads near the chromosome end (reverse 5' end lands in the extension area)
            pos[-1] = L - 2
            reads[c] = {"pos": pos.astype(np.int64), "readlen": ln.astype(np.int64),
                        "is_reverse": rev}
        if tracks is not None and spec["tracks"][i]:
            if spec["tracks"][i] == "empty":
                tracks[c] = (np.zeros(0, np.uint32), np.zeros(0, np.uint32))
                continue
            edges = np.sort(rng.choice(np.arange(0, L), size=2 * max(4, L // 150), replace=False))
            b, e = edges[0::2], edges[1::2]
            e[-1] = L  # interval that ends exactly at the chromosome end
            tracks[c] = (b.astype(np.uint32), e.astype(np.uint32))
    return {"references": refs, "lengths": list(spec["lens"]), "reads": reads, "tracks": tracks}
