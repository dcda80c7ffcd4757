"""BAM and BigWig readers over libpymasc_io.so (include/pymasc_io.h): binary files in, packed
numpy arrays out -- the host-side producers of what the CUDA path consumes.

The reference reads BAM through pysam and BigWig through pyBigWig; neither exists in this image
(SURVEY.md 8f rank 3), so the on-disk formats are decoded natively (pymasc_b200/csrc/pmb_io.cpp:
multi-threaded BGZF inflate, one filtering pass over the records; bbi B+ tree / R-tree walk).
Interfaces follow the reference:

* ``BigWigReader``      PyMaSC/reader/bigwig.pyx:100-200 -- ``chromsizes``, ``fetch(valfilter,
                        chrom)`` (KeyError for an unknown chromosome), ``disable_progress_bar``,
                        ``close``; plus ``fetch_arrays`` -- the packed form the calculator prefers.
* ``BamFile``           the parts of pysam.AlignmentFile that PyMaSC uses (``references``,
                        ``lengths``) plus the packed per-chromosome reads, already filtered by
                        the rules of PyMaSC/handler/read.py:62-90 and :132-149.
* ``estimate_readlen``  PyMaSC/core/readlen.pyx:89-177 (MEAN / MEDIAN / MODE / MIN / MAX over
                        the query lengths of mapped, non-duplicate reads with MAPQ >= criteria).
* ``feed_bam``          the read loop of PyMaSC/handler/calc.py:140-161 as one ``feed_batch``
                        per chromosome.
"""
import ctypes
import os

import numpy as np

from .exceptions import ReadUnsortedError

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PYMASC_IO_LIB") or os.path.join(_HERE, "libpymasc_io.so")

EXPORTS = (
    "pmio_last_error", "pmio_inflate_raw", "pmio_inflate_raw2", "pmio_crc32", "pmio_bam_load", "pmio_bam_free", "pmio_bam_stats_get", "pmio_bam_n_refs",
    "pmio_bam_ref_name", "pmio_bam_ref_len", "pmio_bam_reads", "pmio_bam_packed", "pmio_bam_readlen_hist",
    "pmio_bam_ref_sorted", "pmio_bam_ref_order",
    "pmio_bw_open", "pmio_bw_close", "pmio_bw_n_chroms", "pmio_bw_chrom_name", "pmio_bw_chrom_len",
    "pmio_bw_intervals", "pmio_bw_prefetch",
)
PMIO_ERR_IO, PMIO_ERR_FORMAT, PMIO_ERR_ARG, PMIO_ERR_NOMEM, PMIO_ERR_KEY = -1, -2, -3, -4, -5


class BamStats(ctypes.Structure):
    _fields_ = [("n_records", ctypes.c_uint64), ("n_with_ref", ctypes.c_uint64),
                ("n_unmapped", ctypes.c_uint64), ("n_paired", ctypes.c_uint64),
                ("n_read2", ctypes.c_uint64), ("n_accepted", ctypes.c_uint64),
                ("sorted", ctypes.c_uint32), ("reparsed_shares", ctypes.c_uint32)]


_lib = None


def load_library():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise OSError("libpymasc_io.so is missing (%s). Build it with `python -m pymasc_b200.build`."
                      % LIB_PATH)
    L = ctypes.CDLL(LIB_PATH)
    vp, vpp = ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p)
    u32pp = ctypes.POINTER(ctypes.POINTER(ctypes.c_uint32))
    L.pmio_last_error.restype = ctypes.c_char_p
    L.pmio_inflate_raw.argtypes = [ctypes.c_char_p, ctypes.c_uint64, ctypes.c_char_p, ctypes.c_uint64, ctypes.c_int]
    L.pmio_inflate_raw2.argtypes = [ctypes.c_void_p] * 5
    L.pmio_crc32.argtypes = [ctypes.c_char_p, ctypes.c_uint64, ctypes.c_int]
    L.pmio_crc32.restype = ctypes.c_uint32
    L.pmio_bam_load.argtypes = [ctypes.c_char_p, ctypes.c_int, ctypes.c_int, vpp]
    L.pmio_bam_free.argtypes = [vp]
    L.pmio_bam_free.restype = None
    L.pmio_bam_stats_get.argtypes = [vp, ctypes.POINTER(BamStats)]
    L.pmio_bam_n_refs.argtypes = [vp]
    L.pmio_bam_n_refs.restype = ctypes.c_uint32
    L.pmio_bam_ref_name.argtypes = [vp, ctypes.c_uint32]
    L.pmio_bam_ref_name.restype = ctypes.c_char_p
    L.pmio_bam_ref_len.argtypes = [vp, ctypes.c_uint32]
    L.pmio_bam_ref_len.restype = ctypes.c_uint64
    L.pmio_bam_reads.argtypes = [vp, ctypes.c_uint32, u32pp, u32pp,
                                 ctypes.POINTER(ctypes.POINTER(ctypes.c_uint8)),
                                 ctypes.POINTER(ctypes.c_uint64)]
    u64p = ctypes.POINTER(ctypes.c_uint64)
    L.pmio_bam_packed.argtypes = [vp, ctypes.c_uint32, u32pp, u64p, u64p, u32pp, u64p, u64p,
                                  ctypes.POINTER(ctypes.c_uint32), ctypes.POINTER(ctypes.c_uint32)]
    L.pmio_bam_ref_sorted.argtypes = [vp, ctypes.c_uint32]
    L.pmio_bam_ref_sorted.restype = ctypes.c_int
    L.pmio_bam_ref_order.argtypes = [vp, ctypes.POINTER(ctypes.POINTER(ctypes.c_int32)), u64p, ctypes.POINTER(ctypes.c_int)]
    L.pmio_bam_ref_order.restype = ctypes.c_int
    L.pmio_bam_readlen_hist.argtypes = [vp, u32pp, ctypes.POINTER(ctypes.POINTER(ctypes.c_uint64)),
                                        ctypes.POINTER(ctypes.c_uint32)]
    L.pmio_bw_open.argtypes = [ctypes.c_char_p, vpp]
    L.pmio_bw_close.argtypes = [vp]
    L.pmio_bw_close.restype = None
    L.pmio_bw_n_chroms.argtypes = [vp]
    L.pmio_bw_n_chroms.restype = ctypes.c_uint32
    L.pmio_bw_chrom_name.argtypes = [vp, ctypes.c_uint32]
    L.pmio_bw_chrom_name.restype = ctypes.c_char_p
    L.pmio_bw_chrom_len.argtypes = [vp, ctypes.c_uint32]
    L.pmio_bw_chrom_len.restype = ctypes.c_uint64
    L.pmio_bw_intervals.argtypes = [vp, ctypes.c_char_p, ctypes.c_float, u32pp, u32pp,
                                    ctypes.POINTER(ctypes.POINTER(ctypes.c_float)),
                                    ctypes.POINTER(ctypes.c_uint64)]
    L.pmio_bw_prefetch.argtypes = [vp, ctypes.c_float, ctypes.c_int]
    _lib = L
    return L


def inflate_raw(data, out_len, use_zlib=False):
    """One raw-DEFLATE stream -> bytes of exactly `out_len` (ValueError otherwise), through the
    library's own decoder or through zlib."""
    L = load_library()
    out = ctypes.create_string_buffer(max(int(out_len), 1))
    rc = L.pmio_inflate_raw(bytes(data), len(data), out, int(out_len), int(bool(use_zlib)))
    if rc != 0:
        _raise(L, rc)
    return out.raw[:int(out_len)]


def inflate_raw2(data_a, out_len_a, data_b, out_len_b):
    """Two raw-DEFLATE streams through the library's two-stream decoder: a pair of bytes objects,
    None where a stream was refused."""
    L = load_library()
    datas = (bytes(data_a), bytes(data_b))
    lens = (int(out_len_a), int(out_len_b))
    outs = [ctypes.create_string_buffer(max(n, 1)) for n in lens]
    ins = (ctypes.c_char_p * 2)(*datas)
    in_len = (ctypes.c_uint64 * 2)(len(datas[0]), len(datas[1]))
    out_ptr = (ctypes.c_void_p * 2)(ctypes.addressof(outs[0]), ctypes.addressof(outs[1]))
    out_len = (ctypes.c_uint64 * 2)(*lens)
    ok = (ctypes.c_int * 2)(0, 0)
    rc = L.pmio_inflate_raw2(ins, in_len, out_ptr, out_len, ok)
    if rc != 0:
        _raise(L, rc)
    return tuple(outs[i].raw[:lens[i]] if ok[i] else None for i in range(2))


def crc32(data, use_zlib=False):
    """CRC-32 of a bytes-like object through the library (the BGZF block check)."""
    data = bytes(data)
    return int(load_library().pmio_crc32(data, len(data), int(bool(use_zlib))))


def _raise(L, rc, path=None):
    msg = L.pmio_last_error().decode(errors="replace")
    if rc == PMIO_ERR_KEY:
        raise KeyError(msg)
    if rc == PMIO_ERR_IO:
        raise IOError(msg)
    if rc == PMIO_ERR_NOMEM:
        raise MemoryError(msg)
    raise ValueError(msg)


def _copy(ptr, n, dtype):
    if n == 0:
        return np.zeros(0, dtype)
    return np.ctypeslib.as_array(ptr, shape=(n,)).astype(dtype, copy=True)


# ---------------------------------------------------------------------------------------- BAM
class _LazyReads(object):
    """chrom -> {pos, readlen, is_reverse}; a chromosome's arrays are copied out of the native
    handle on first access (most callers only ever need the packed form)."""

    def __init__(self, bam):
        self._bam, self._got = bam, {}

    def __getitem__(self, chrom):
        if chrom not in self._got:
            self._got[chrom] = self._bam._reads_of(chrom)
        return self._got[chrom]

    def __contains__(self, chrom):
        return chrom in self._bam._counts

    def __iter__(self):
        return iter(self._bam._counts)

    def __len__(self):
        return len(self._bam._counts)

    def keys(self):
        return self._bam._counts.keys()

    def items(self):
        return ((c, self[c]) for c in self._bam._counts)


class BamFile(object):
    """A decoded BAM file: header plus the accepted reads per reference.

    ``reads[chrom]`` is a dict of ``pos`` (1-based leftmost base), ``readlen`` (query length) and
    ``is_reverse`` in file order -- the arguments of ``CCBitArrayCalculator.feed_batch``;
    ``packed(chrom)`` is the same chromosome already reduced by the calculator's feed rules --
    the arguments of ``feed_packed``.
    Filter: skip read2, MAPQ < mapq_criteria, unmapped, duplicate, records without a reference or
    a CIGAR (PyMaSC/handler/read.py:62-90, :132-143)."""

    def __init__(self, path, mapq_criteria=1, threads=0):
        path = str(path)
        self._h = None
        if not os.path.exists(path):
            raise IOError("input file '{0}' dose not exist.".format(path))
        L = self._L = load_library()
        h = ctypes.c_void_p()
        rc = L.pmio_bam_load(path.encode(), int(mapq_criteria), int(threads), ctypes.byref(h))
        if rc != 0:
            _raise(L, rc)
        self._h = h
        n = L.pmio_bam_n_refs(h)
        self.references = tuple(L.pmio_bam_ref_name(h, i).decode("utf-8", "surrogateescape") for i in range(n))
        self.lengths = tuple(int(L.pmio_bam_ref_len(h, i)) for i in range(n))
        self._index = {name: i for i, name in enumerate(self.references)}
        st = BamStats()
        L.pmio_bam_stats_get(h, ctypes.byref(st))
        self.stats = {k: int(getattr(st, k)) for k, _ in BamStats._fields_}
        self.sorted = bool(st.sorted)
        self._counts = {}                       # chromosomes with accepted reads, header order
        cnt = ctypes.c_uint64()
        for i, name in enumerate(self.references):
            L.pmio_bam_reads(h, i, None, None, None, ctypes.byref(cnt))
            if cnt.value:
                self._counts[name] = int(cnt.value)
        # order as the reference sees it (per chromosome, and the sequence of chromosomes in the file)
        self.ref_sorted = {name: bool(L.pmio_bam_ref_sorted(h, i)) for i, name in enumerate(self.references)}
        ro, rn, rov = ctypes.POINTER(ctypes.c_int32)(), ctypes.c_uint64(), ctypes.c_int()
        L.pmio_bam_ref_order(h, ctypes.byref(ro), ctypes.byref(rn), ctypes.byref(rov))
        self.ref_order = [self.references[i] for i in _copy(ro, rn.value, np.int64).tolist()]
        self.ref_order_overflow = bool(rov.value)
        self.reads = _LazyReads(self)
        hl, hc, hn = ctypes.POINTER(ctypes.c_uint32)(), ctypes.POINTER(ctypes.c_uint64)(), ctypes.c_uint32()
        L.pmio_bam_readlen_hist(h, ctypes.byref(hl), ctypes.byref(hc), ctypes.byref(hn))
        self.readlen_counter = dict(zip(_copy(hl, hn.value, np.int64).tolist(),
                                        _copy(hc, hn.value, np.int64).tolist()))
        self.path = path
        self.mapq_criteria = int(mapq_criteria)

    def _reads_of(self, chrom):
        if chrom not in self._counts:
            raise KeyError(chrom)
        p, q = ctypes.POINTER(ctypes.c_uint32)(), ctypes.POINTER(ctypes.c_uint32)()
        r, cnt = ctypes.POINTER(ctypes.c_uint8)(), ctypes.c_uint64()
        self._L.pmio_bam_reads(self._h, self._index[chrom], ctypes.byref(p), ctypes.byref(q),
                               ctypes.byref(r), ctypes.byref(cnt))
        return {"pos": _copy(p, cnt.value, np.int64), "readlen": _copy(q, cnt.value, np.int64),
                "is_reverse": _copy(r, cnt.value, bool)}

    def packed(self, chrom):
        """The chromosome's reads after the feed rules of mscc.pyx:370-418: dict of ``fwd``
        (sorted unique forward 5' positions, uint32), ``fwd_len_sum``, ``rev`` (sorted unique
        reverse 5' ends), ``rev_len_sum``, ``first_pos``, ``last_pos``.  The chromosome's own
        reads must be in coordinate order."""
        if chrom not in self._counts:
            raise KeyError(chrom)
        if not self.ref_sorted[chrom]:
            raise ReadUnsortedError
        f, r = ctypes.POINTER(ctypes.c_uint32)(), ctypes.POINTER(ctypes.c_uint32)()
        nf, nr, fs, rs = ctypes.c_uint64(), ctypes.c_uint64(), ctypes.c_uint64(), ctypes.c_uint64()
        p0, p1 = ctypes.c_uint32(), ctypes.c_uint32()
        rc = self._L.pmio_bam_packed(self._h, self._index[chrom], ctypes.byref(f), ctypes.byref(nf),
                                     ctypes.byref(fs), ctypes.byref(r), ctypes.byref(nr), ctypes.byref(rs),
                                     ctypes.byref(p0), ctypes.byref(p1))
        if rc != 0:
            _raise(self._L, rc)
        return {"fwd": _copy(f, nf.value, np.uint32), "fwd_len_sum": int(fs.value),
                "rev": _copy(r, nr.value, np.uint32), "rev_len_sum": int(rs.value),
                "first_pos": int(p0.value), "last_pos": int(p1.value)}

    def close(self):
        if getattr(self, "_h", None):
            self._L.pmio_bam_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:       # noqa: BLE001 -- interpreter shutdown
            pass


def _counter_arrays(counter):
    lengths = np.array(sorted(counter), dtype=np.int64)
    return lengths, np.array([counter[int(l)] for l in lengths], dtype=np.int64)


def _estimate(counter, esttype):
    """The five estimators of core/readlen.pyx:22-86 over a {query length: reads} counter.
    MEAN rounds the weighted mean (Python ``round``); MEDIAN is the lower-half boundary of the
    cumulative counts, the rounded midpoint of two lengths when an even total splits exactly
    between them; MODE is the most frequent length, ties broken like the reference: by the order in
    which the lengths first occur in the file (the counter is built in that order)."""
    lengths, counts = _counter_arrays(counter)
    if esttype == "MIN":
        return int(lengths[0])
    if esttype == "MAX":
        return int(lengths[-1])
    if esttype == "MODE":       # readlen.pyx:81, literally: a stable sort by count, ties in first-seen order
        return int([k for k, v in sorted(counter.items(), key=lambda x: x[1])][-1])
    total = int(counts.sum())
    if esttype == "MEAN":
        return int(round(int((lengths * counts).sum()) / float(total)))
    if esttype != "MEDIAN":
        raise KeyError(esttype)
    cum = np.cumsum(counts)
    half = total / 2
    at = int(np.searchsorted(cum, half, side="left"))          # first length whose cumulative count >= half
    if total % 2 == 0 and cum[at] == half:
        return int(round((int(lengths[at]) + float(lengths[at + 1])) / 2))
    return int(lengths[at])


ESTFUNCTIONS = {name: (lambda counter, _n=name: _estimate(counter, _n))
                for name in ("MEAN", "MEDIAN", "MODE", "MIN", "MAX")}


def estimate_readlen(path, esttype="MEDIAN", mapq_criteria=1, threads=0):
    """core/readlen.pyx:89-177.  `path` may be a BamFile already decoded with the same MAPQ."""
    bam = path if hasattr(path, "readlen_counter") else BamFile(path, mapq_criteria, threads)
    if bam.mapq_criteria != int(mapq_criteria):
        raise ValueError("BamFile was decoded with another MAPQ criterion")
    if not bam.readlen_counter:
        raise ValueError("There is no read to estimate the read length from.")
    return ESTFUNCTIONS[esttype](bam.readlen_counter)


def feed_bam(calc, bam, only=None):
    """Feed every accepted read of `bam` to a calculator, one ``feed_packed`` per chromosome in
    the order the chromosomes appear in the file (handler/calc.py:140-153).
    `only`: a set of chromosome names; reads elsewhere are skipped BEFORE any order check, as the
    reference does (handler/calc.py:143-144 skips, then ``_check_pos``, mscc.pyx:351-366): a file
    that is unsorted only inside an excluded chromosome feeds fine; a selected chromosome whose
    reads go backwards, or that is left for another selected chromosome and revisited, raises
    ReadUnsortedError."""
    if bam.ref_order_overflow:
        raise ReadUnsortedError
    order = []
    for chrom in bam.ref_order:
        if only is not None and chrom not in only:
            continue
        if order and order[-1] == chrom:
            continue                    # only excluded chromosomes lay in between
        if chrom in order or not bam.ref_sorted[chrom]:
            raise ReadUnsortedError
        order.append(chrom)
    # the per-read rules (duplicates, read-length sums) were applied natively while decoding
    for chrom in order:
        calc.feed_packed(chrom, **bam.packed(chrom))


# ------------------------------------------------------------------------------------- BigWig
class BigWigReader(object):
    """reader/bigwig.pyx:100-200 over the native decoder."""

    def __init__(self, path):
        path_str = str(path)
        if not os.path.exists(path_str) and path_str != '-':
            raise IOError("input file '{0}' dose not exist.".format(path_str))
        if path_str == '-':
            raise ValueError("BigWig input from stdin is not supported")
        self.path = path_str
        self.closed = False
        self._L = load_library()
        h = ctypes.c_void_p()
        rc = self._L.pmio_bw_open(path_str.encode(), ctypes.byref(h))
        if rc != 0:
            self.closed = True
            if rc == PMIO_ERR_FORMAT:
                raise IOError("Failed to open BigWig file: {} ({})".format(
                    path_str, self._L.pmio_last_error().decode()))
            _raise(self._L, rc)
        self._h = h
        n = self._L.pmio_bw_n_chroms(h)
        self.chromsizes = {self._L.pmio_bw_chrom_name(h, i).decode("utf-8", "surrogateescape"): int(self._L.pmio_bw_chrom_len(h, i))
                           for i in range(n)}

    def fetch_arrays(self, valfilter, chrom, with_values=False):
        """(begins, ends[, values]) of the intervals with value >= valfilter: 0-based half-open
        uint32 arrays in file order.  KeyError for a chromosome the file does not have."""
        if self.closed:
            raise ValueError("I/O operation on closed BigWig file")
        if chrom not in self.chromsizes:
            raise KeyError(chrom)
        b, e = ctypes.POINTER(ctypes.c_uint32)(), ctypes.POINTER(ctypes.c_uint32)()
        v, n = ctypes.POINTER(ctypes.c_float)(), ctypes.c_uint64()
        rc = self._L.pmio_bw_intervals(self._h, chrom.encode("utf-8", "surrogateescape"), float(valfilter), ctypes.byref(b),
                                       ctypes.byref(e), ctypes.byref(v), ctypes.byref(n))
        if rc != 0:
            _raise(self._L, rc)
        out = (_copy(b, n.value, np.uint32), _copy(e, n.value, np.uint32))
        return out + (_copy(v, n.value, np.float32),) if with_values else out

    def prefetch(self, valfilter, threads=0):
        """Decode every chromosome at `valfilter` now, on `threads` threads (0 = all cores);
        fetch / fetch_arrays with that threshold then only copy."""
        if self.closed:
            raise ValueError("I/O operation on closed BigWig file")
        rc = self._L.pmio_bw_prefetch(self._h, float(valfilter), int(threads))
        if rc != 0:
            _raise(self._L, rc)

    def fetch(self, valfilter, chrom):
        """List of (begin, end, value) tuples, as BigWigFileIterator yields them."""
        b, e, v = self.fetch_arrays(valfilter, chrom, with_values=True)
        return list(zip(b.tolist(), e.tolist(), v.astype(np.float64).tolist()))

    def disable_progress_bar(self):
        pass

    def close(self):
        if not self.closed:
            self._L.pmio_bw_close(self._h)
            self._h = None
            self.closed = True

    def __del__(self):
        try:
            self.close()
        except Exception:       # noqa: BLE001 -- interpreter shutdown
            pass
