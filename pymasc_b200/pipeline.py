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
import logging
import time

from . import io_native
from .calculator import CCBitArrayCalculator
from .multifile import MultiFileCalculation

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


class CachedTrack(object):
    """A BigWigReader whose per-chromosome interval arrays are decoded once and reused by every
    file of a multi-file run (the reference re-reads the BigWig for each file)."""

    def __init__(self, reader):
        self._reader = reader
        self.chromsizes = reader.chromsizes
        self._cache = {}

    def fetch_arrays(self, valfilter, chrom):
        key = (float(valfilter), chrom)
        if key not in self._cache:
            self._cache[key] = self._reader.fetch_arrays(valfilter, chrom)     # KeyError passes through
        return self._cache[key]

    def fetch(self, valfilter, chrom):
        return self._reader.fetch(valfilter, chrom)

    def disable_progress_bar(self):
        pass

    def close(self):
        self._reader.close()


def run_files(bam_paths, bigwig_path=None, max_shift=1000, mapq_criteria=1, read_len=None,
              esttype="MEDIAN", skip_ncc=False, threads=0, engine=None):
    """Cross-correlation results of one or more BAM files.

    Returns (results, info): ``results[i]`` is the genome-wide result of ``bam_paths[i]``
    (NCCGenomeWideResult / MSCCGenomeWideResult / BothGenomeWideResult, as
    ``CCBitArrayCalculator.get_whole_result`` of the reference returns them); ``info`` holds
    per-file read lengths, read statistics and wall-clock seconds per stage."""
    if isinstance(bam_paths, (str, bytes)) or hasattr(bam_paths, "__fspath__"):
        bam_paths = [bam_paths]
    info = {"files": [], "seconds": {"decode": 0.0, "feed": 0.0, "correlate": 0.0}}
    feeder = CachedTrack(io_native.BigWigReader(bigwig_path)) if bigwig_path is not None else None
    calcs = []
    try:
        for path in bam_paths:
            t0 = time.perf_counter()
            bam = io_native.BamFile(path, mapq_criteria, threads)
            t1 = time.perf_counter()
            rl = read_len
            if rl is None:                       # handler/calc.py:85-98
                rl = io_native.estimate_readlen(bam, esttype, mapq_criteria)
            if rl > max_shift:
                raise ValueError("Read length ({}) seems to be longer than shift size ({})."
                                 .format(rl, max_shift))
            lengths = list(bam.lengths)
            if feeder is not None:
                lengths = reconcile_chromosome_sizes(bam.references, lengths, feeder.chromsizes)
            calc = CCBitArrayCalculator(max_shift, rl, list(bam.references), lengths, feeder,
                                        skip_ncc, engine=engine)
            io_native.feed_bam(calc, bam)
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
