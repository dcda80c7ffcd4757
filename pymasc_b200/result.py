"""Result containers handed back by the B200 calculator.

Drop-in rule: when the PyMaSC package is importable (the calculator is running inside
PyMaSC, under PyMaSC/handler), the reference's OWN classes from ``PyMaSC.result`` are
returned so that ``stats.py``, ``output/*`` and ``aggregate_results`` see exactly the
types they test with ``isinstance`` (PyMaSC/stats.py:611-642, PyMaSC/result.py:301-356).
Stand-alone (bench, GPU box, tests) the mirrors below are used: same class names, same
field names, same ``calc_cc`` arithmetic in the same operation order, so the float64
curves are bit-identical given identical integer counts.

Mirrored from (behaviour, not code):
  NCCResult / calc_cc ............ PyMaSC/result.py:68-89    denom = L - k, means = sum / L
  MSCCResult / calc_cc ........... PyMaSC/result.py:92-118   totlen[k] = mappable_len[|rl-1-k|]
  _calc_cc ....................... PyMaSC/result.py:42-65    all-zero ccbins -> all-NaN
  Empty*Result.create_empty ...... PyMaSC/result.py:144-228
  *GenomeWideResult .............. PyMaSC/result.py:270-298
  FP error policy ................ PyMaSC/utils/calc.py:152-169 (raise, then retry ignoring)
"""
import os
from dataclasses import dataclass, field
from typing import Any, Dict, Optional

import numpy as np

def _normalised_cc(forward_sum, reverse_sum, ccbins, totlen, denom):
    bins = np.array(ccbins, dtype=np.int64)
    if bins.sum() == 0:
        return np.full_like(bins, np.nan, dtype=np.float64)

    def body():
        fmean = forward_sum / totlen
        rmean = reverse_sum / totlen
        fvar = fmean * (1 - fmean)
        rvar = rmean * (1 - rmean)
        prod = fmean * rmean
        geo = (fvar * rvar) ** 0.5
        return (bins / denom - prod) / geo

    try:
        with np.errstate(divide="raise", invalid="raise"):
            return body()
    except (FloatingPointError, ZeroDivisionError):
        with np.errstate(divide="ignore", invalid="ignore"):
            return body()



# ---------------------------------------------------------------------------- cc from the count vectors
# SURVEY.md 8(f) row 4 (`calc_cc`, PyMaSC/result.py:42-118): the calculator holds the per-shift
# counts as int64 numpy vectors when they come back from the device, while the result classes
# keep them as Python lists (the reference's field type) and `calc_cc()` converts every list
# back to an array -- four conversions of max_shift+1 Python ints per chromosome, most of the
# host time of a finished batch.  These two functions compute the same float64 curve straight
# from the vectors, in calc_cc's own operation order (bit-identical: tests/test_calculator_host.py),
# so the calculator assigns `result.cc` instead of calling `result.calc_cc()`.  It stays on the
# host: max_shift+1 float64 values per chromosome give a device nothing to do.
def ncc_cc(forward_sum, reverse_sum, ccbins, genomelen, max_shift):
    """NCCResult.calc_cc (PyMaSC/result.py:82-89) from an int64 vector."""
    shifts = np.arange(max_shift + 1, dtype=np.float64)
    return _normalised_cc(float(forward_sum), float(reverse_sum), ccbins[:max_shift + 1], genomelen,
                          genomelen - shifts)


def mappable_len_array(dmp, read_len, max_shift):
    """float64 form of the reference's mappable_len list (mscc.pyx:271, :292-298; unset slots NaN,
    as np.array(list_with_None, dtype=float64) gives them)."""
    S = max_shift + 1
    n = min(read_len, S)
    ml = np.full(read_len, np.nan)
    ml[read_len - n:read_len] = dmp[:n][::-1]
    return np.concatenate((ml, np.asarray(dmp[2 * read_len - 1:S], dtype=np.float64)))


def mscc_cc(forward_sum, reverse_sum, ccbins, dmp, read_len, max_shift):
    """MSCCResult.calc_cc (PyMaSC/result.py:104-118) from int64 vectors; `dmp` = popc(D_k) by shift."""
    n = max_shift + 1
    ml = mappable_len_array(dmp, read_len, max_shift)
    tot = np.concatenate((ml[:read_len][::-1], ml[1:]))[:n]
    return _normalised_cc(np.asarray(forward_sum[:n], dtype=np.float64), np.asarray(reverse_sum[:n], dtype=np.float64),
                          ccbins[:n], tot, tot)


try:  # inside PyMaSC: hand back the reference's own types
    if os.environ.get("PYMASC_B200_MIRROR_RESULTS"):  # tests: force the mirrors
        raise ImportError("mirrors forced")
    from PyMaSC.result import (  # type: ignore  # noqa: F401
        NCCResult, MSCCResult, BothChromResult,
        NCCGenomeWideResult, MSCCGenomeWideResult, BothGenomeWideResult,
        EmptyNCCResult, EmptyMSCCResult,
    )
    USING_REFERENCE_TYPES = True
except Exception:  # stand-alone mirrors
    USING_REFERENCE_TYPES = False

    @dataclass
    class NCCResult:
        max_shift: int
        read_len: int
        genomelen: int
        forward_sum: int
        reverse_sum: int
        forward_read_len_sum: int
        reverse_read_len_sum: int
        ccbins: Any
        cc: Any = field(init=False, default=None)

        def calc_cc(self):
            shifts = np.array(range(self.max_shift + 1), dtype=np.float64)
            self.cc = _normalised_cc(float(self.forward_sum), float(self.reverse_sum),
                                     self.ccbins[:self.max_shift + 1], self.genomelen,
                                     self.genomelen - shifts)

    @dataclass
    class MSCCResult:
        max_shift: int
        read_len: int
        genomelen: int
        forward_sum: Any
        reverse_sum: Any
        forward_read_len_sum: Optional[int]
        reverse_read_len_sum: Optional[int]
        ccbins: Any
        mappable_len: Any
        cc: Any = field(init=False, default=None)

        def calc_cc(self):
            assert self.mappable_len is not None
            ml = np.array(self.mappable_len, dtype=np.float64)
            # index |read_len-1-k|: the first read_len entries mirrored, then the tail
            tot = np.concatenate((ml[:self.read_len][::-1], ml[1:]))[:self.max_shift + 1]
            n = self.max_shift + 1
            self.cc = _normalised_cc(np.array(self.forward_sum[:n], dtype=np.float64),
                                     np.array(self.reverse_sum[:n], dtype=np.float64),
                                     self.ccbins[:n], tot, tot)

    @dataclass
    class BothChromResult:
        chrom: Optional[NCCResult]
        mappable_chrom: Optional[MSCCResult]

    @dataclass
    class EmptyNCCResult(NCCResult):
        @classmethod
        def create_empty(cls, genome_length, max_shift, read_len):
            r = cls(max_shift=max_shift, read_len=read_len, genomelen=genome_length,
                    forward_sum=0, reverse_sum=0, forward_read_len_sum=0,
                    reverse_read_len_sum=0, ccbins=[0.0] * (max_shift + 1))
            r.calc_cc()
            return r

    @dataclass
    class EmptyMSCCResult(MSCCResult):
        @classmethod
        def create_empty(cls, genome_length, max_shift, read_len):
            n = max_shift + 1
            r = cls(max_shift=max_shift, read_len=read_len, genomelen=genome_length,
                    forward_sum=np.zeros(n, dtype=np.int64),
                    reverse_sum=np.zeros(n, dtype=np.int64),
                    forward_read_len_sum=0, reverse_read_len_sum=0,
                    ccbins=[0.0] * n, mappable_len=tuple([0] * n))
            r.calc_cc()
            return r

    @dataclass
    class NCCGenomeWideResult:
        genomelen: int
        forward_sum: int
        reverse_sum: int
        chroms: Dict[str, NCCResult]
        forward_read_len_sum: int
        reverse_read_len_sum: int

    @dataclass
    class MSCCGenomeWideResult:
        genomelen: int
        chroms: Dict[str, MSCCResult]
        forward_read_len_sum: int
        reverse_read_len_sum: int

    @dataclass
    class BothGenomeWideResult:
        genomelen: int
        forward_sum: int
        reverse_sum: int
        chroms: Dict[str, NCCResult]
        mappable_chroms: Dict[str, MSCCResult]
        forward_read_len_sum: int
        reverse_read_len_sum: int
