"""Checker engine for host-logic tests: answers calculator jobs with the CPU oracle.

Lives in tests/ on purpose -- the product package has no CPU path.  It lets the `not gpu`
suite exercise CCBitArrayCalculator's bookkeeping (feed rules, result types, list lengths)
against the reference's vectors without a device.
"""
from oracle import cc_oracle
from pymasc_b200 import _shim


class OracleEngine(object):
    def __init__(self):
        self.batches = []

    def run(self, jobs):
        self.batches.append(len(jobs))
        outs = []
        track = None
        for j in jobs:
            if j.flags & _shim.FLAG_SAME_TRACK:      # the previous job's intervals (pymasc_b200.h)
                assert track is not None and j.map_begins.size == 0
            elif j.flags & _shim.FLAG_MSCC:
                track = (j.map_begins, j.map_ends)
            else:
                track = None
            if j.flags & _shim.FLAG_AUTOCORR:
                outs.append({"dmp": cc_oracle.mappable_autocorr(
                    j.chrom_len, j.read_len, j.max_shift, j.map_begins, j.map_ends,
                    lo=j.lo, hi=j.hi if j.hi else None)})
                continue
            mscc = bool(j.flags & _shim.FLAG_MSCC)
            outs.append(cc_oracle.chrom_counts(
                j.chrom_len, j.read_len, j.max_shift, j.fwd_pos, j.rev_pos,
                track[0] if mscc else None, track[1] if mscc else None,
                skip_ncc=not (j.flags & _shim.FLAG_NCC),
                lo=j.lo, hi=j.hi if j.hi else None))
        return outs
