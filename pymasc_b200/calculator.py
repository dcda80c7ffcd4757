"""B200 drop-in for PyMaSC's bit-array calculator.

``CCBitArrayCalculator`` here has the constructor, attributes, methods, error behaviour and
result objects of the reference class of the same name
(PyMaSC/core/bitarray/mscc.pyx:41-483; interface PyMaSC/interfaces/calculator.py:15-124), so it
can be constructed at PyMaSC/handler/factory.py:247 in its place and everything downstream
(result.py, stats.py, output/*) runs unchanged.

What differs is where the work happens.  The reference sets one bit per ``feed_*`` call in a
host BitArray and, at every chromosome switch, runs (max_shift+1) x ~10 whole-array passes on
one CPU core.  Here ``feed_*`` only appends the 5' position to a packed uint32 buffer and keeps
the reference's host-side bookkeeping (sort check, duplicate rules, read-length sums); closed
chromosomes are queued and handed to the device in ONE batch -- bit-array build and the
shift-window kernel run in libpymasc_b200.so (pymasc_b200/csrc) -- when results are needed
(``flush``, ``finishup_calculation``, ``get_result``, ``get_whole_result``).

There is no CPU fallback: without the CUDA library and an sm_100 device the first batch raises
``DeviceError``.
"""
import logging
import os
from array import array

import numpy as np

from . import _shim
from .exceptions import PositionOutOfRange, ReadUnsortedError
from .result import mscc_cc, ncc_cc
from .result import (
    BothChromResult, BothGenomeWideResult, EmptyMSCCResult, EmptyNCCResult, MSCCGenomeWideResult,
    MSCCResult, NCCGenomeWideResult, NCCResult,
)

logger = logging.getLogger(__name__)


def _ints(v):
    """Python ints of a count vector (ndarray.tolist converts at C speed; the reference's
    result fields are plain lists of ints)."""
    return v.tolist() if hasattr(v, "tolist") else [int(x) for x in v]


class DeviceEngine(object):
    """Runs batches of chromosome jobs on one B200 through the C ABI."""

    def __init__(self, device=None):
        self._device = device
        self._ordinal = None if device is None or isinstance(device, _shim.Device) else int(device)
        if not isinstance(device, _shim.Device):
            self._device = None

    @property
    def device(self):
        if self._device is None:      # opened lazily: spawned workers pick their own GPU
            self._device = _shim.Device(self._ordinal)
        return self._device

    def run(self, jobs):
        return self.device.run_batch(jobs)


class _NoProgress(object):
    """progress_bar=None: the reference creates its own terminal ProgressBar here
    (mscc.pyx:140); per-shift ticks do not exist on the device, so nothing is drawn."""

    def __getattr__(self, name):
        return lambda *a, **k: None


class _Pending(object):
    __slots__ = ("chrom", "fwd", "rev", "intervals", "fwd_len_sum", "rev_len_sum")


class CCBitArrayCalculator(object):
    """Naive and mappability-sensitive strand cross-correlation counts on a B200.

    Args (same order and meaning as mscc.pyx:97-98):
        max_shift, read_len, references, lengths, bwfeeder=None, skip_ncc=False,
        logger_lock=None, progress_bar=None
    Extra keyword (not in the reference): ``engine`` -- object with ``run(jobs)``; defaults to
    ``DeviceEngine()``.  Tests inject a checker engine here; the product never does.
    """

    def __init__(self, max_shift, read_len, references, lengths, bwfeeder=None, skip_ncc=False,
                 logger_lock=None, progress_bar=None, engine=None):
        self.MAPPABILITY_THRESHOLD = 1.         # mscc.pyx:116
        self.EXTRA_ALLOCATE_SIZE = 100          # mscc.pyx:118
        self._max_shift = int(max_shift)
        self._read_len = int(read_len)
        self.references = references
        self.ref2genomelen = dict(zip(references, lengths))
        self.genomelen = sum(lengths)
        self.ref2ncc_result = {}
        self.ref2mscc_result = {}
        self.forward_sum = self.reverse_sum = 0
        self.forward_read_len_sum = self.reverse_read_len_sum = 0

        self._chr = ''
        self._solved_chr = []
        self._buff_flashed = False
        self._array_extend_size = self._read_len + self._max_shift + self.EXTRA_ALLOCATE_SIZE

        self._bwfeeder = bwfeeder
        self._skip_ncc = bool(skip_ncc)
        self._logger_lock = logger_lock
        # mscc.pyx:140-142: a caller's hook (a ProgressHook posting to the report queue in `-p N`
        # workers, utils/progress.py:173-200) gets the mappability bar of every chromosome and no
        # per-shift bar ("bitarray progress stacks in multiprocessing")
        self._progress = progress_bar if progress_bar else _NoProgress()
        self._allow_bitarray_progress = progress_bar is None
        self._engine = engine if engine is not None else DeviceEngine()

        self._pending = []                      # closed chromosomes awaiting the device
        self._fwd = self._rev_end = self._rev_len = None
        self._nbits = 0
        self._last_pos = self._last_forward_pos = 0
        self._forward_read_len_sum = 0
        self._packed = None                     # feed_packed: (fwd, fwd_len_sum, rev, rev_len_sum)

    # readonly attributes of the reference (mscc.pyx:75-79)
    max_shift = property(lambda self: self._max_shift)
    read_len = property(lambda self: self._read_len)
    skip_ncc = property(lambda self: self._skip_ncc)
    logger_lock = property(lambda self: self._logger_lock)

    # ------------------------------------------------------------------ logging
    def _logging_info(self, msg):
        if self._logger_lock:
            self._logger_lock.acquire()
        logger.info(msg)
        if self._logger_lock:
            self._logger_lock.release()

    # ------------------------------------------------------------------ progress (mscc.pyx:151-159)
    def _init_mappability_progress(self, chrom):
        self._progress.enable_bar()
        self._progress.reset_progress("@@@@@^" * 12, '^', ' ', chrom, self.ref2genomelen[chrom])

    def _init_bitarray_progress(self, chrom):
        if self._allow_bitarray_progress:
            self._progress.reset_progress("xxxxxX" * 12, 'X', ' ', chrom, self._max_shift)
        else:
            self._progress.disable_bar()

    # ------------------------------------------------------------------ feeding (host)
    def _init_buff(self):
        # mscc.pyx:161-171 -- the bit arrays become packed position buffers
        self._nbits = self.ref2genomelen[self._chr] + self._array_extend_size
        self._fwd = array('I')
        self._rev_end = array('I')
        self._rev_len = array('I')
        self._last_pos = 0
        self._last_forward_pos = 0
        self._forward_read_len_sum = 0
        self._packed = None

    def _check_pos(self, chrom, pos):
        # mscc.pyx:351-366
        if chrom != self._chr:
            if self._chr != '':
                if chrom in self._solved_chr:
                    raise ReadUnsortedError
                self._solved_chr.append(self._chr)
                self._close_chromosome()
                self._buff_flashed = False
            self._chr = chrom
            self._init_buff()
            self._logging_info("Loading {} reads to bit array...".format(chrom))
        elif self._packed is not None:
            raise ValueError("chromosome %s was fed packed (feed_packed); it takes no further reads" % chrom)
        if pos < self._last_pos:
            raise ReadUnsortedError
        self._last_pos = pos

    def feed_forward_read(self, chrom, pos, readlen):
        """pos: 1-based leftmost base.  mscc.pyx:370-393."""
        self._check_pos(chrom, pos)
        if self._last_forward_pos == pos:
            return None                         # duplicated read
        self._last_forward_pos = pos
        if not 0 <= pos < self._nbits:
            raise PositionOutOfRange("forward read at %s:%d beyond the bit array" % (chrom, pos))
        self._forward_read_len_sum += readlen
        self._fwd.append(pos)

    def feed_reverse_read(self, chrom, pos, readlen):
        """The bit set is pos + readlen - 1, the 5' end of the reverse read.  mscc.pyx:397-418."""
        self._check_pos(chrom, pos)
        end = pos + readlen - 1
        if not 0 <= end < self._nbits:
            raise PositionOutOfRange("reverse read end at %s:%d beyond the bit array" % (chrom, end))
        # first-occurrence rule for the read-length sum is applied when the chromosome closes
        self._rev_end.append(end)
        self._rev_len.append(readlen)

    def feed_batch(self, chrom, pos, readlen, is_reverse):
        """Vectorised feed of one chromosome's reads in coordinate order (numpy arrays).
        Equivalent to calling feed_forward_read / feed_reverse_read per read."""
        pos = np.asarray(pos, dtype=np.int64)
        readlen = np.asarray(readlen, dtype=np.int64)
        is_reverse = np.asarray(is_reverse, dtype=bool)
        if pos.size == 0:
            return
        self._check_pos(chrom, int(pos[0]))
        if pos.size > 1 and np.any(pos[1:] < pos[:-1]):
            raise ReadUnsortedError
        self._last_pos = int(pos[-1])
        fpos, flen = pos[~is_reverse], readlen[~is_reverse]
        if fpos.size:
            keep = np.ones(fpos.size, dtype=bool)
            keep[1:] = fpos[1:] != fpos[:-1]
            keep[0] = fpos[0] != self._last_forward_pos
            fpos, flen = fpos[keep], flen[keep]
            if fpos.size:
                if fpos[0] < 0 or fpos[-1] >= self._nbits:
                    raise PositionOutOfRange("forward read on %s beyond the bit array" % chrom)
                self._last_forward_pos = int(fpos[-1])
                self._forward_read_len_sum += int(flen.sum())
                self._fwd.frombytes(fpos.astype(np.uint32).tobytes())
        rend = pos[is_reverse] + readlen[is_reverse] - 1
        if rend.size:
            if rend.min() < 0 or rend.max() >= self._nbits:
                raise PositionOutOfRange("reverse read end on %s beyond the bit array" % chrom)
            self._rev_end.frombytes(rend.astype(np.uint32).tobytes())
            self._rev_len.frombytes(readlen[is_reverse].astype(np.uint32).tobytes())

    def feed_packed(self, chrom, fwd, fwd_len_sum, rev, rev_len_sum, first_pos, last_pos):
        """A WHOLE chromosome's reads already reduced by the feed rules (what a native decoder
        produces, pymasc_b200.io_native.BamFile.packed): ``fwd`` sorted unique forward 5'
        positions with ``fwd_len_sum`` the query lengths of the reads kept by the duplicate rule
        (mscc.pyx:388-393), ``rev`` sorted unique reverse 5' ends with ``rev_len_sum`` over the
        first read of each end (:416-418); ``first_pos`` / ``last_pos`` are the leftmost
        positions of the first and last read fed.  Equivalent to feeding those reads one by one;
        the chromosome cannot take further reads afterwards."""
        if chrom == self._chr:
            raise ValueError("feed_packed needs a chromosome that has not been fed yet")
        self._check_pos(chrom, int(first_pos))
        fwd = np.ascontiguousarray(fwd, dtype=np.uint32)
        rev = np.ascontiguousarray(rev, dtype=np.uint32)
        if fwd.size and int(fwd[-1]) >= self._nbits:
            raise PositionOutOfRange("forward read on %s beyond the bit array" % chrom)
        if rev.size and int(rev[-1]) >= self._nbits:
            raise PositionOutOfRange("reverse read end on %s beyond the bit array" % chrom)
        self._last_pos = int(last_pos)
        self._last_forward_pos = int(fwd[-1]) if fwd.size else 0
        self._packed = (fwd, int(fwd_len_sum), rev, int(rev_len_sum))

    # ------------------------------------------------------------------ mappability (host)
    def _load_intervals(self, chrom):
        """(begins, ends) uint32 arrays of the intervals with value >= 1.0, or None without a
        feeder.  KeyError when the track lacks the chromosome.  mscc.pyx:327-349."""
        if not self._bwfeeder:
            return None
        if hasattr(self._bwfeeder, "fetch_arrays"):          # packed fast path
            b, e = self._bwfeeder.fetch_arrays(self.MAPPABILITY_THRESHOLD, chrom)
        else:
            feeder = self._bwfeeder.fetch(self.MAPPABILITY_THRESHOLD, chrom)
            rows = [(iv[0], iv[1]) for iv in feeder]
            arr = np.asarray(rows, dtype=np.int64).reshape(-1, 2)
            b, e = arr[:, 0], arr[:, 1]
        self._logging_info("Loading {} mappability to bit array...".format(chrom))
        # mscc.pyx:342-347: the reference ticks once per interval while it fills the bit array;
        # here the intervals arrive as arrays, so the bar goes from empty to full in one update
        self._init_mappability_progress(chrom)
        self._progress.update(self.ref2genomelen[chrom])
        self._progress.clean()
        nbits = self.ref2genomelen[chrom] + self._array_extend_size
        if (isinstance(b, np.ndarray) and isinstance(e, np.ndarray) and b.dtype == np.uint32 and e.dtype == np.uint32
                and b.flags.c_contiguous and e.flags.c_contiguous):     # the native readers' arrays: no conversions
            if e.size and int(e.max()) >= nbits:
                raise PositionOutOfRange("mappability interval on %s beyond the bit array" % chrom)
            return b, e
        b = np.ascontiguousarray(b, dtype=np.int64)
        e = np.ascontiguousarray(e, dtype=np.int64)
        if b.size and (b.min() < 0 or e.max() >= nbits):
            # the reference aborts on a BitArray assert here (bit_array.c:825)
            raise PositionOutOfRange("mappability interval on %s beyond the bit array" % chrom)
        return b.astype(np.uint32), e.astype(np.uint32)

    # ------------------------------------------------------------------ closing a chromosome
    def _close_chromosome(self):
        """The host half of _calc_correlation (mscc.pyx:217-282): settle the read-length sums,
        reserve the result slots in the reference's insertion order, queue the device job."""
        if self._chr == '' or self._buff_flashed:
            return
        chrom = self._chr
        p = _Pending()
        p.chrom = chrom
        if self._packed is not None:
            p.fwd, p.fwd_len_sum, p.rev, p.rev_len_sum = self._packed
            self._packed = None
            self._queue(p)
            return
        p.fwd = np.frombuffer(self._fwd, dtype=np.uint32).copy() if len(self._fwd) else np.zeros(0, np.uint32)
        rev_end = np.frombuffer(self._rev_end, dtype=np.uint32) if len(self._rev_end) else np.zeros(0, np.uint32)
        rev_len = np.frombuffer(self._rev_len, dtype=np.uint32) if len(self._rev_len) else np.zeros(0, np.uint32)
        # reverse duplicate rule (mscc.pyx:416-418): only the first read to set a bit counts
        uniq, first = np.unique(rev_end, return_index=True)
        p.rev = uniq.astype(np.uint32)
        p.fwd_len_sum = int(self._forward_read_len_sum)
        p.rev_len_sum = int(rev_len[first].sum()) if first.size else 0
        self._queue(p)

    def _queue(self, p):
        chrom = p.chrom
        self.forward_read_len_sum += p.fwd_len_sum
        self.reverse_read_len_sum += p.rev_len_sum
        if not self._skip_ncc:
            self.ref2ncc_result[chrom] = None          # reserved; filled by _compute_pending
        try:
            p.intervals = self._load_intervals(chrom)
        except KeyError as e:
            self._logging_info("Mappability for '{}' not found. "
                               "Skip calc mappability sensitive CC.".format(e.args[0]))
            p.intervals = None
        if p.intervals is not None:
            self.ref2mscc_result[chrom] = None
        self._logging_info("Calculate cross-correlation for {}...".format(chrom))
        if not self._skip_ncc or p.intervals is not None:
            self._pending.append(p)
        self._fwd = self._rev_end = self._rev_len = None

    def _job_of(self, p):
        flags = (0 if self._skip_ncc else _shim.FLAG_NCC) | \
                (_shim.FLAG_MSCC if p.intervals is not None else 0)
        b, e = p.intervals if p.intervals is not None else (None, None)
        return _shim.Job(self.ref2genomelen[p.chrom], self._read_len, self._max_shift, flags,
                         p.fwd, p.rev, b, e)

    def _mappable_len_list(self, dmp):
        # mscc.pyx:271, :292-298 -- popc(D_k) re-indexed by |read_len-1-k|: shifts k < read_len
        # fill ml[read_len-1-k], shifts k >= 2*read_len-1 are appended, the ones between repeat
        # values already stored and are dropped
        rl, S = self._read_len, self._max_shift + 1
        d = dmp.tolist() if hasattr(dmp, "tolist") else [int(x) for x in dmp]
        n = min(rl, S)
        ml = [None] * rl
        ml[rl - n:rl] = d[:n][::-1]
        return ml + d[2 * rl - 1:S]

    def _store(self, p, out):
        chrom, L = p.chrom, self.ref2genomelen[p.chrom]
        # mscc.pyx:286, :317, :325: the per-shift bar -- one tick per finished chromosome here
        self._init_bitarray_progress(chrom)
        self._progress.update(self._max_shift)
        if not self._skip_ncc:
            self.forward_sum += out["fwd_total"]
            self.reverse_sum += out["rev_total"]
            r = NCCResult(max_shift=self._max_shift, read_len=self._read_len, genomelen=L,
                          forward_sum=out["fwd_total"], reverse_sum=out["rev_total"],
                          forward_read_len_sum=p.fwd_len_sum, reverse_read_len_sum=p.rev_len_sum,
                          ccbins=_ints(out["ncc"]))
            # calc_cc()'s curve, computed from the vector itself (result.ncc_cc: bit-identical)
            r.cc = ncc_cc(out["fwd_total"], out["rev_total"], np.asarray(out["ncc"]), L, self._max_shift)
            self.ref2ncc_result[chrom] = r
        if p.intervals is not None:
            r = MSCCResult(max_shift=self._max_shift, read_len=self._read_len, genomelen=L,
                           forward_sum=_ints(out["mfwd"]),
                           reverse_sum=_ints(out["mrev"]),
                           forward_read_len_sum=p.fwd_len_sum, reverse_read_len_sum=p.rev_len_sum,
                           ccbins=_ints(out["mscc"]),
                           mappable_len=self._mappable_len_list(out["dmp"]))
            r.cc = mscc_cc(np.asarray(out["mfwd"]), np.asarray(out["mrev"]), np.asarray(out["mscc"]),
                           np.asarray(out["dmp"]), self._read_len, self._max_shift)
            self.ref2mscc_result[chrom] = r
        self._progress.clean()

    def _track_key(self, chrom):
        """Names the mappability track of a job for pymasc_b200.multifile: readers of the same
        BigWig file (the reference opens one per input file, reader/bigwig.pyx:129-137) share a
        key through their ``path``; anything else is keyed by object identity."""
        path = getattr(self._bwfeeder, "path", None)
        if isinstance(path, (str, os.PathLike)):
            return ("file", os.path.abspath(os.fspath(path)), float(self.MAPPABILITY_THRESHOLD), chrom)
        return ("object", id(self._bwfeeder), chrom)

    def _collect_jobs(self, fill=()):
        """Device jobs of every queued chromosome plus the mappable-length autocorrelation of
        the read-less chromosomes in `fill` (mscc.pyx:181-215).  Returns (jobs, fills); the
        outputs of `jobs`, in order, go to _apply_results."""
        jobs = [self._job_of(p) for p in self._pending]
        for j, p in zip(jobs, self._pending):
            if p.intervals is not None:
                j.track_key = self._track_key(p.chrom)
        fills = []
        for chrom in fill:
            self._chr = chrom                                   # side effect of _fill_result
            if chrom not in self.ref2ncc_result:
                self.ref2ncc_result[chrom] = EmptyNCCResult.create_empty(
                    self.ref2genomelen[chrom], self._max_shift, self._read_len)
            if not self._bwfeeder or chrom in self.ref2mscc_result:
                continue
            res = self.ref2mscc_result[chrom] = EmptyMSCCResult.create_empty(
                self.ref2genomelen[chrom], self._max_shift, self._read_len)
            try:
                iv = self._load_intervals(chrom)
            except KeyError:
                continue
            self._logging_info("Calc {} mappable length...".format(chrom))
            job = _shim.Job(self.ref2genomelen[chrom], self._read_len, self._max_shift,
                            _shim.FLAG_AUTOCORR, None, None, iv[0], iv[1])
            job.track_key = self._track_key(chrom)
            jobs.append(job)
            fills.append(res)
        return jobs, fills

    def _apply_results(self, outs, fills):
        for p, out in zip(self._pending, outs):
            self._store(p, out)
        for res, out in zip(fills, outs[len(self._pending):]):
            res.mappable_len = tuple(_ints(out["dmp"]))
        self._pending = []

    def _compute_pending(self, fill=()):
        """One device batch for everything _collect_jobs returns."""
        jobs, fills = self._collect_jobs(fill)
        if not jobs:
            return
        self._apply_results(self._engine.run(jobs), fills)

    # ------------------------------------------------------------------ public protocol
    def flush(self, chrom=None):
        """mscc.pyx:173-179."""
        if self._chr != '' and not self._buff_flashed:
            self._close_chromosome()
        self._compute_pending(fill=() if chrom is None else (chrom,))
        self._buff_flashed = True

    def finishup_calculation(self):
        """mscc.pyx:420-439: flush the last chromosome, then fill every reference."""
        if self._chr != '' and not self._buff_flashed:
            self._close_chromosome()
        last = self._chr
        self._compute_pending(fill=[last] + list(self.references))
        self._buff_flashed = True

    def get_result(self, chrom):
        """mscc.pyx:441-447."""
        if chrom not in self.ref2ncc_result and chrom not in self.ref2mscc_result:
            raise KeyError(chrom)
        self._compute_pending()
        return BothChromResult(chrom=self.ref2ncc_result.get(chrom),
                               mappable_chrom=self.ref2mscc_result.get(chrom))

    def get_whole_result(self):
        """mscc.pyx:449-483."""
        self._compute_pending()
        if not self.ref2mscc_result:
            assert self.ref2ncc_result, "No results available for either NCC or MSCC."
            return NCCGenomeWideResult(
                genomelen=self.genomelen, forward_sum=self.forward_sum, reverse_sum=self.reverse_sum,
                chroms=self.ref2ncc_result.copy(),
                forward_read_len_sum=self.forward_read_len_sum,
                reverse_read_len_sum=self.reverse_read_len_sum)
        elif not self.ref2ncc_result:
            return MSCCGenomeWideResult(
                genomelen=self.genomelen, chroms=self.ref2mscc_result.copy(),
                forward_read_len_sum=self.forward_read_len_sum,
                reverse_read_len_sum=self.reverse_read_len_sum)
        return BothGenomeWideResult(
            genomelen=self.genomelen, forward_sum=self.forward_sum, reverse_sum=self.reverse_sum,
            chroms=self.ref2ncc_result.copy(), mappable_chroms=self.ref2mscc_result.copy(),
            forward_read_len_sum=self.forward_read_len_sum,
            reverse_read_len_sum=self.reverse_read_len_sum)
