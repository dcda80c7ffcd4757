"""ctypes front end of oracle/cc_oracle.c -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
``--impl reference`` legs may import this module; the product package
``pymasc_b200`` never does.  Parity status: pinned (see cc_oracle.c header).

`chrom_counts` produces, for one chromosome, the same integers the reference's
``CCBitArrayCalculator._calc_correlation`` (PyMaSC/core/bitarray/mscc.pyx:217-325)
appends to its result lists, keyed by shift k.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "libcc_oracle.so")
_lib = None

EXTRA_ALLOCATE_SIZE = 100  # mscc.pyx:118


def build(force=False):
    """Compile cc_oracle.c with gcc (make -C oracle)."""
    if force or not os.path.exists(_LIB_PATH) or (
            os.path.getmtime(_LIB_PATH) < os.path.getmtime(os.path.join(_HERE, "cc_oracle.c"))):
        subprocess.check_call(["make", "-s", "-C", _HERE, "-B", "all"])
    return _LIB_PATH


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        L = ctypes.CDLL(_LIB_PATH)
        u64p = ctypes.POINTER(ctypes.c_uint64)
        u32p = ctypes.POINTER(ctypes.c_uint32)
        i64p = ctypes.POINTER(ctypes.c_int64)
        L.orc_set_bits.argtypes = [u64p, ctypes.c_uint64, u32p, ctypes.c_uint64]
        L.orc_set_bits.restype = ctypes.c_int
        L.orc_fill_intervals.argtypes = [u64p, ctypes.c_uint64, u32p, u32p, ctypes.c_uint64]
        L.orc_fill_intervals.restype = ctypes.c_int
        L.orc_count.argtypes = [u64p, ctypes.c_uint64]
        L.orc_count.restype = ctypes.c_uint64
        L.orc_count_range.argtypes = [u64p, ctypes.c_uint64, ctypes.c_uint64, ctypes.c_uint64]
        L.orc_count_range.restype = ctypes.c_uint64
        L.orc_cc_counts.argtypes = [u64p, u64p, u64p, ctypes.c_uint64, ctypes.c_int64,
                                    ctypes.c_int64, ctypes.c_uint64, ctypes.c_uint64,
                                    i64p, i64p, i64p, i64p, i64p]
        L.orc_cc_counts.restype = ctypes.c_int
        L.orc_mappable_autocorr.argtypes = [u64p, ctypes.c_uint64, ctypes.c_int64, i64p]
        L.orc_mappable_autocorr.restype = ctypes.c_int
        _lib = L
    return _lib


def _p(a, ct):
    return a.ctypes.data_as(ctypes.POINTER(ct)) if a is not None else None


def array_bits(chrom_len, read_len, max_shift):
    """Length of the reference's per-chromosome bit arrays (mscc.pyx:134, :165)."""
    return int(chrom_len) + int(read_len) + int(max_shift) + EXTRA_ALLOCATE_SIZE


def make_bits(nbits, positions):
    """Zeroed array + bit_array_set_bit per position.  Raises IndexError where the
    reference would abort on its assert."""
    w = np.zeros((nbits + 63) // 64, dtype=np.uint64)
    pos = np.ascontiguousarray(positions, dtype=np.uint32)
    if lib().orc_set_bits(_p(w, ctypes.c_uint64), nbits, _p(pos, ctypes.c_uint32), pos.size):
        raise IndexError("position beyond bit array")
    return w


def make_mappability(nbits, begins, ends):
    w = np.zeros((nbits + 63) // 64, dtype=np.uint64)
    b = np.ascontiguousarray(begins, dtype=np.uint32)
    e = np.ascontiguousarray(ends, dtype=np.uint32)
    if lib().orc_fill_intervals(_p(w, ctypes.c_uint64), nbits, _p(b, ctypes.c_uint32),
                                _p(e, ctypes.c_uint32), b.size):
        raise IndexError("interval beyond bit array")
    return w


def chrom_counts(chrom_len, read_len, max_shift, fwd_pos, rev_pos,
                 map_begins=None, map_ends=None, skip_ncc=False, lo=0, hi=None):
    """Per-shift counts for one chromosome (or the position shard [lo, hi)).

    fwd_pos: 1-based forward 5' ends; rev_pos: 1-based reverse 5' ends
    (= pos + readlen - 1, mscc.pyx:416-417).  Duplicates collapse (bit set twice).
    Returns dict with int64 arrays of max_shift+1 entries: ncc, dmp, mfwd, mrev,
    mscc (MSCC keys only with mappability) and ints fwd_total, rev_total.
    """
    L = lib()
    nbits = array_bits(chrom_len, read_len, max_shift)
    if hi is None:
        hi = nbits
    F = make_bits(nbits, fwd_pos)
    R = make_bits(nbits, rev_pos)
    M = None
    if map_begins is not None:
        M = make_mappability(nbits, map_begins, map_ends)
    S = max_shift + 1
    out = {}
    ncc = None if skip_ncc else np.zeros(S, dtype=np.int64)
    dmp = mf = mr = ms = None
    if M is not None:
        dmp, mf, mr, ms = (np.zeros(S, dtype=np.int64) for _ in range(4))
    rc = L.orc_cc_counts(_p(F, ctypes.c_uint64), _p(R, ctypes.c_uint64), _p(M, ctypes.c_uint64),
                         nbits, read_len, max_shift, lo, hi,
                         _p(ncc, ctypes.c_int64), _p(dmp, ctypes.c_int64), _p(mf, ctypes.c_int64),
                         _p(mr, ctypes.c_int64), _p(ms, ctypes.c_int64))
    assert rc == 0
    if ncc is not None:
        out["ncc"] = ncc
    if M is not None:
        out.update(dmp=dmp, mfwd=mf, mrev=mr, mscc=ms)
    out["fwd_total"] = int(L.orc_count_range(_p(F, ctypes.c_uint64), nbits, lo, hi))
    out["rev_total"] = int(L.orc_count_range(_p(R, ctypes.c_uint64), nbits, lo, hi))
    return out


def mappable_autocorr(chrom_len, read_len, max_shift, map_begins, map_ends, lo=0, hi=None):
    """mappable_len tuple of a read-less chromosome (mscc.pyx:207-215); with lo/hi the
    partial sums over j in [lo, hi) of a position shard (they add up to the whole)."""
    nbits = array_bits(chrom_len, read_len, max_shift)
    M = make_mappability(nbits, map_begins, map_ends)
    out = np.zeros(max_shift + 1, dtype=np.int64)
    if lo == 0 and (hi is None or hi >= nbits):
        lib().orc_mappable_autocorr(_p(M, ctypes.c_uint64), nbits, max_shift, _p(out, ctypes.c_int64))
    else:   # sum_j M[j] M[j+k] is the NCC count of F = R = M
        rc = lib().orc_cc_counts(_p(M, ctypes.c_uint64), _p(M, ctypes.c_uint64), None, nbits,
                                 read_len, max_shift, lo, min(hi, nbits), _p(out, ctypes.c_int64),
                                 None, None, None, None)
        assert rc == 0
    return out


def mappable_len_list(dmp, read_len, max_shift):
    """Re-index popc(D_k) by |read_len-1-k| exactly as mscc.pyx:271,:292-298 does:
    list starts as [None]*read_len, k < read_len writes index read_len-1-k,
    read_len <= k < 2*read_len-1 writes nothing, later k append."""
    ml = [None] * read_len
    for k in range(max_shift + 1):
        if k < read_len:
            ml[read_len - k - 1] = int(dmp[k])
        elif k < read_len * 2 - 1:
            pass
        else:
            ml.append(int(dmp[k]))
    return ml
