"""From files to genome-wide results: BAM (+ BigWig) -> packed arrays -> B200 -> result objects.

This is the reference's per-file workflow (PyMaSC/handler/calc.py:60-161: read-length check,
chromosome-size reconciliation with the BigWig, calculator construction, the read loop,
``finishup_calculation``) with the native readers of ``io_native`` in place of pysam/pyBigWig
and one ``feed_batch`` per chromosome in place of one Python call per read.  Several BAM files
over one BigWig are finished together (``MultiFileCalculation``): the mappability bit arrays
are built once per chromosome for all files.

It stops where the hot path stops: what comes back are the reference's result objects
(``get_whole_result()``), ready for the reference's own stats.py / output code.
"""
import fnmatch
import logging
import time
from itertools import groupby

from . import io_native, io_text
from .calculator import CCBitArrayCalculator
from .multifile import MultiFileCalculation

MAPPABILITY_THRESHOLD = 1.      # what every CCBitArrayCalculator asks the track for (mscc.pyx:116)

logger = logging.getLogger(__name__)


def reconcile_chromosome_sizes(references, lengths, external_sizes):
    """PyMaSC/reader/bam.py:217-255: where the BAM header and the BigWig disagree on a
    chromosome's length the longer one is used."""
    out = []
    for ref, bam_size in zip(references, lengths):
        ext = external_sizes.get(ref)
        if ext is not None and ext != bam_size:
            logger.warning("'%s' reference length mismatch: SAM/BAM -> %d, External -> %d",
                           ref, bam_size, ext)
            out.append(max(bam_size, ext))
        else:
            out.append(bam_size)
    return out


class SamFile(object):
    """A text SAM file behind the interface ``run_files`` needs from ``io_native.BamFile``
    (pysam reads both formats for the reference; the text twin is parsed in Python,
    pymasc_b200/io_text.py, and is meant for small files and tests)."""

    def __init__(self, path, mapq_criteria=1):
        refs, lens, self.reads, self.readlen_counter = io_text.read_sam(path, mapq_criteria, with_readlen_counter=True)
        self.references, self.lengths = tuple(refs), tuple(lens)
        self.mapq_criteria = int(mapq_criteria)
        self.stats = {"n_accepted": sum(len(r["pos"]) for r in self.reads.values())}
        self.path = str(path)

    def close(self):
        pass


def open_alignment(path, mapq_criteria=1, threads=0):
    """BamFile for BGZF input, SamFile for text (decided by the gzip magic, as htslib does)."""
    with open(path, "rb") as fh:
        magic = fh.read(2)
    if magic == b"\x1f\x8b":
        return io_native.BamFile(path, mapq_criteria, threads)
    return SamFile(path, mapq_criteria)


class NoTargetChromosomes(ValueError):
    """The chromosome filter leaves nothing (reader/bam.py:201-203, BAMNoTargetChroms)."""


def filter_chroms(chroms, filters):
    """The set of chromosome names that -i / --include-chrom and -e / --exclude-chrom select
    (utils/calc.py:49-91).  `filters` is the reference's ``chromfilter`` list, in command-line
    order: (True, [patterns]) for an include option, (False, [patterns]) for an exclude option,
    shell-style patterns.  Runs of the same kind act as one option.  An exclude keeps what it
    does not match and hands what it DID match to the option after it, so that a later include
    can take names back; an include narrows the names still under consideration to its matches;
    what the last option leaves is selected if that option is an include."""
    if filters is None:
        return set(chroms)
    considered, selected, including = set(chroms), set(), False
    for including, group in groupby(filters, key=lambda f: f[0]):
        patterns = {p for _, pats in group for p in pats}
        matched = {c for p in patterns for c in fnmatch.filter(considered, p)}
        if not including:
            selected |= considered - matched
        considered = matched
    if including:
        selected |= considered
    return selected


def apply_chromfilter(references, lengths, chromfilter):
    """(references, lengths) of the selected chromosomes, header order (reader/bam.py:180-215)."""
    keep = filter_chroms(references, chromfilter)
    if not keep:
        raise NoTargetChromosomes("There is no targeted chromosomes.")
    pairs = [(r, n) for r, n in zip(references, lengths) if r in keep]
    return [r for r, _ in pairs], [n for _, n in pairs]


def feed_alignment(calc, aln, only=None):
    """Every accepted read to the calculator, file order; reads on chromosomes outside `only`
    (the filtered reference list) are skipped as handler/calc.py:143-144 skips them."""
    if isinstance(aln, io_native.BamFile):
        return io_native.feed_bam(calc, aln, only)
    for chrom, r in aln.reads.items():          # file order; feed_batch raises ReadUnsortedError itself
        if only is None or chrom in only:
            calc.feed_batch(chrom, r["pos"], r["readlen"], r["is_reverse"])


class CachedTrack(object):
    """A BigWigReader whose per-chromosome interval arrays are decoded once and reused by every
    file of a multi-file run (the reference re-reads the BigWig for each file)."""

    def __init__(self, reader):
        self._reader = reader
        self.path = reader.path
        self.chromsizes = reader.chromsizes
        self._cache = {}

    def fetch_arrays(self, valfilter, chrom):
        key = (float(valfilter), chrom)
        if key not in self._cache:
            self._cache[key] = self._reader.fetch_arrays(valfilter, chrom)     # KeyError passes through
        return self._cache[key]

    def prefetch(self, valfilter, threads=0):
        """Every chromosome decoded now, in parallel (pmio_bw_prefetch)."""
        self._reader.prefetch(valfilter, threads)

    def fetch(self, valfilter, chrom):
        return self._reader.fetch(valfilter, chrom)

    def disable_progress_bar(self):
        pass

    def close(self):
        self._reader.close()


def precalc_mappability(bigwig_path, max_shift=1000, max_readlen=1000, stats_path=None, threads=0, engine=None):
    """The body of ``pymasc-precalc`` (calcmappablelen.py:54-58): mappable-length statistics of a
    BigWig track for shifts up to what (max_shift, max_readlen) require, written to the JSON
    cache that ``pymasc`` reads (next to the BigWig unless `stats_path` says otherwise).  Every
    chromosome in one device batch.  Returns the handler (``mappable_len``,
    ``chrom2mappable_len``, ``map_path``)."""
    from .mappability import MappabilityHandler
    if stats_path is not None and str(stats_path) == str(bigwig_path):      # calcmappablelen.py:51-52
        stats_path = None
    handler = MappabilityHandler(bigwig_path, max_shift, max_readlen, stats_path, threads or 1, engine=engine)
    try:
        handler.calc_mappability()
        handler.save_mappability_stats()
    finally:
        handler.close()
    return handler


def run_files(bam_paths, bigwig_path=None, max_shift=1000, mapq_criteria=1, read_len=None,
              esttype="MEDIAN", skip_ncc=False, threads=0, engine=None, chromfilter=None):
    """Cross-correlation results of one or more BAM files.

    Returns (results, info): ``results[i]`` is the genome-wide result of ``bam_paths[i]``
    (NCCGenomeWideResult / MSCCGenomeWideResult / BothGenomeWideResult, as
    ``CCBitArrayCalculator.get_whole_result`` of the reference returns them); ``info`` holds
    per-file read lengths, read statistics and wall-clock seconds per stage.  ``chromfilter`` is
    the reference's include / exclude list (see filter_chroms); the read-length estimate looks at
    every chromosome, as core/readlen.pyx does."""
    if isinstance(bam_paths, (str, bytes)) or hasattr(bam_paths, "__fspath__"):
        bam_paths = [bam_paths]
    info = {"files": [], "seconds": {"decode": 0.0, "feed": 0.0, "correlate": 0.0}}
    feeder = CachedTrack(io_native.BigWigReader(bigwig_path)) if bigwig_path is not None else None
    calcs = []
    try:
        if feeder is not None:                   # one parallel pass instead of a decode per chromosome
            t0 = time.perf_counter()
            feeder.prefetch(MAPPABILITY_THRESHOLD, threads)
            info["seconds"]["decode"] += time.perf_counter() - t0
        for path in bam_paths:
            t0 = time.perf_counter()
            bam = open_alignment(path, mapq_criteria, threads)
            t1 = time.perf_counter()
            rl = read_len
            if rl is None:                       # handler/calc.py:85-98
                rl = io_native.estimate_readlen(bam, esttype, mapq_criteria)
            if rl > max_shift:
                raise ValueError("Read length ({}) seems to be longer than shift size ({})."
                                 .format(rl, max_shift))
            references, lengths = apply_chromfilter(list(bam.references), list(bam.lengths), chromfilter)
            if feeder is not None:
                lengths = reconcile_chromosome_sizes(references, lengths, feeder.chromsizes)
            calc = CCBitArrayCalculator(max_shift, rl, references, lengths, feeder, skip_ncc, engine=engine)
            feed_alignment(calc, bam, None if chromfilter is None else set(references))
            t2 = time.perf_counter()
            info["files"].append({"path": str(path), "read_len": int(rl), "stats": bam.stats})
            info["seconds"]["decode"] += t1 - t0
            info["seconds"]["feed"] += t2 - t1
            calcs.append(calc)
            bam.close()                          # the packed arrays were copied out; free the decoder's memory now
        t0 = time.perf_counter()
        MultiFileCalculation(calcs, engine=engine).finishup_calculation()
        results = [c.get_whole_result() for c in calcs]
        info["seconds"]["correlate"] = time.perf_counter() - t0
    finally:
        if feeder is not None:
            feeder.close()
    info["seconds"]["total"] = sum(info["seconds"].values())
    return results, info
