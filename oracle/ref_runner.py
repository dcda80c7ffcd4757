"""Drive the compiled, unmodified reference calculator (oracle/_ref) from packed arrays.

TEST INFRASTRUCTURE ONLY: imported by tests/, tests/golden/make_golden.py and
bench.py's cpu_baseline / ``--impl reference`` legs.  Never by pymasc_b200.

The reference is fed exactly as PyMaSC/handler/read.py:144-147 feeds it (one
``feed_forward_read`` / ``feed_reverse_read`` call per read, in coordinate order) and
finished as PyMaSC/handler/calc.py:159-161 (single process) or
PyMaSC/handler/worker.py:224-234 (``-p N`` worker) does.
"""
import glob
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def _popcnt(popcnt):
    """None: the process-wide choice, PYMASC_REF_POPCNT=1 (one interpreter can import only one
    build of the package, so bench.py picks the variant per spawned worker through the env)."""
    return os.environ.get("PYMASC_REF_POPCNT") == "1" if popcnt is None else bool(popcnt)


def ref_dir(popcnt=None):
    return os.path.join(HERE, "_ref_popcnt" if _popcnt(popcnt) else "_ref")


def available(popcnt=None):
    return bool(glob.glob(os.path.join(ref_dir(popcnt), "PyMaSC", "core", "bitarray", "mscc*.so")))


def import_reference(popcnt=None):
    """Put oracle/_ref (and the pyBigWig stand-in) on sys.path; return the calculator class."""
    if not available(popcnt):
        raise ImportError("compiled reference missing: run python oracle/build_ref.py")
    for p in (os.path.join(HERE, "ref_stubs"), ref_dir(popcnt)):
        if p not in sys.path:
            sys.path.insert(0, p)
    from PyMaSC.core.bitarray.mscc import CCBitArrayCalculator
    return CCBitArrayCalculator


class ArrayFeeder(object):
    """Duck-typed ``bwfeeder``: ``fetch(thr, chrom)`` must return the reference's own
    BigWigFileIterator (typed at mscc.pyx:329,:336) and raise KeyError for an unknown
    chromosome (reader/bigwig.pyx:166-167)."""

    def __init__(self, tracks):
        # tracks: {chrom: (begins, ends)} -- all intervals have value 1.0
        import_reference()
        from PyMaSC.reader.bigwig import BigWigFileIterator
        self._it = BigWigFileIterator
        self._tracks = tracks

    def fetch(self, valfilter, chrom):
        if chrom not in self._tracks:
            raise KeyError(chrom)
        b, e = self._tracks[chrom]
        return self._it([(int(x), int(y), 1.0) for x, y in zip(b, e)])

    def disable_progress_bar(self):
        pass

    def close(self):
        pass


class _Quiet(object):
    """progress_bar argument: silences the reference's terminal bars."""

    def __getattr__(self, name):
        return lambda *a, **k: None


def merged_feed_order(fwd_pos, fwd_len, rev_pos, rev_len):
    """Coordinate-sorted interleaving of the two strands (stable, forward first on ties),
    as a BAM iterator would deliver them.  rev_pos is the read's leftmost base."""
    pos = np.concatenate([np.asarray(fwd_pos, np.int64), np.asarray(rev_pos, np.int64)])
    ln = np.concatenate([np.asarray(fwd_len, np.int64), np.asarray(rev_len, np.int64)])
    rev = np.concatenate([np.zeros(len(fwd_pos), bool), np.ones(len(rev_pos), bool)])
    order = np.argsort(pos, kind="stable")
    return pos[order], ln[order], rev[order]


def run_reference(max_shift, read_len, references, lengths, reads_by_chrom,
                  tracks=None, skip_ncc=False, popcnt=False, finish=True):
    """reads_by_chrom: ordered {chrom: (pos, readlen, is_reverse)} arrays, pos = 1-based
    leftmost base in feed order.  Returns the calculator after finishup_calculation()."""
    cls = import_reference(popcnt)
    feeder = ArrayFeeder(tracks) if tracks is not None else None
    calc = cls(max_shift, read_len, list(references), list(lengths), feeder, skip_ncc,
               None, _Quiet())
    for chrom, (pos, ln, rev) in reads_by_chrom.items():
        ff, fr = calc.feed_forward_read, calc.feed_reverse_read
        for p, l, r in zip(pos.tolist(), ln.tolist(), rev.tolist()):
            if r:
                fr(chrom, p, l)
            else:
                ff(chrom, p, l)
    if finish:
        calc.finishup_calculation()
    return calc
