"""Run in a fresh interpreter by tests/test_reference_seam.py: the reference's OWN factory
(PyMaSC/handler/factory.py from oracle/_ref, unmodified) with the one-line swap of
INTEGRATION.md section 1 applied -- the name `CCBitArrayCalculator` in that module now is the B200
class -- builds the calculator through `create_calculator(config)` (factory.py:108-136,
:224-258: the positional constructor call at :247-256), for the three calculation targets.
The reference's fixture reads are fed read by read as handler/read.py:144-147 does, the
calculator is finished as handler/calc.py:159-161 does, and a snapshot of every result is
printed as JSON together with the calls the progress hook received.

argv[1]: "oracle" (CPU checker engine, no GPU) or "device"; argv[2]: scratch directory.
"""
import json
import os
import shutil
import sys
from pathlib import Path

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "oracle", "ref_stubs"), os.path.join(ROOT, "oracle", "_ref"),
          os.path.join(ROOT, "tests"), ROOT):
    sys.path.insert(0, p)

import PyMaSC.handler.factory as factory              # noqa: E402  the reference, unmodified
from PyMaSC.interfaces.config import (                # noqa: E402
    Algorithm, CalculationTarget, EstimationType, PyMaSCConfig)
import PyMaSC.result as ref_result                    # noqa: E402
import helpers                                        # noqa: E402
import pymasc_b200                                    # noqa: E402
from pymasc_b200 import io_native                     # noqa: E402

mode, scratch = sys.argv[1], sys.argv[2]

if mode == "oracle":
    from oracle_engine import OracleEngine

    class Swapped(pymasc_b200.CCBitArrayCalculator):
        """The drop-in with the CPU checker engine behind it (tests only: the factory's
        positional call has no way to pass `engine`)."""

        def __init__(self, *args, **kwargs):
            kwargs["engine"] = OracleEngine()
            super(Swapped, self).__init__(*args, **kwargs)
else:
    Swapped = pymasc_b200.CCBitArrayCalculator

factory.CCBitArrayCalculator = Swapped                # the swap at factory.py:27

# The factory opens the track with the reference's BigWigReader (reader/bigwig.pyx), which needs
# pyBigWig; oracle/ref_stubs/pyBigWig.py serves X.bigwig from a bedGraph twin beside it.  Write
# that twin (and the chromosome sizes) from the binary fixture with this repo's native reader.
bw_path = os.path.join(scratch, "hg19_36mer-test.bigwig")
shutil.copyfile(os.path.join(helpers.GOLDEN, "hg19_36mer-test.bigwig"), bw_path)
native = io_native.BigWigReader(bw_path)
with open(os.path.join(scratch, "hg19_36mer-test.bedGraph"), "w") as fh:
    for chrom in native.chromsizes:
        b, e = native.fetch_arrays(1.0, chrom)
        for x, y in zip(b.tolist(), e.tolist()):
            fh.write("%s\t%d\t%d\t1\n" % (chrom, x, y))
with open(os.path.join(scratch, "hg19.chrom.sizes"), "w") as fh:
    for chrom, size in native.chromsizes.items():
        fh.write("%s\t%d\n" % (chrom, size))
native.close()


class RecordingHook(object):
    """Stands where a ProgressHook would (utils/progress.py:173-200)."""

    def __init__(self):
        self.calls = []

    def __getattr__(self, name):
        def record(*args):
            self.calls.append([name] + [a for a in args])
        return record


fx = helpers.load_fixture()
out = {}
for label, target, with_track in (("fixture_ncc", CalculationTarget.NCC, False),
                                  ("fixture_both", CalculationTarget.BOTH, True),
                                  ("fixture_mscc_only", CalculationTarget.MSCC, True)):
    config = PyMaSCConfig(
        max_shift=300, mapq_criteria=10, target=target, implementation=Algorithm.BITARRAY, nproc=1,
        esttype=EstimationType.MEDIAN, chi2_pval=0.05, mv_avr_filter_len=15, filter_mask_len=5,
        min_calc_width=50, read_length=36, ref2lengths=dict(zip(fx["references"], fx["lengths"])),
        mappability_path=Path(bw_path) if with_track else None)
    hook = RecordingHook()
    calc, bwfeeder = factory.create_calculator(config, None, hook)
    # tests/unit/test_factory_architecture.py:216-217 of the reference
    assert isinstance(calc, factory.CCBitArrayCalculator)
    assert (bwfeeder is not None) == with_track
    assert (calc.max_shift, calc.read_len, calc.skip_ncc, calc.logger_lock) == \
        (300, 36, target is CalculationTarget.MSCC, None)
    for pos, readlen, is_rev in zip(fx["pos"].tolist(), fx["readlen"].tolist(), fx["is_reverse"].tolist()):
        if is_rev:
            calc.feed_reverse_read("chr1", pos, readlen)
        else:
            calc.feed_forward_read("chr1", pos, readlen)
    calc.finishup_calculation()
    whole = calc.get_whole_result()
    assert type(whole).__module__ == ref_result.__name__          # the reference's own result classes
    chrom = calc.get_result("chr1")
    assert isinstance(chrom, ref_result.BothChromResult)
    # the calculator assigns .cc from the count vectors; the reference's own calc_cc() on the same
    # objects must give the same float64 values bit for bit (PyMaSC/result.py:82-118)
    import numpy as np
    for res in list(whole.chroms.values()) + list(getattr(whole, "mappable_chroms", {}).values()):
        cc0 = np.array(res.cc, dtype=np.float64, copy=True)
        res.calc_cc()
        assert np.array_equal(cc0, np.asarray(res.cc, dtype=np.float64), equal_nan=True), type(res).__name__
    if bwfeeder is not None:
        bwfeeder.close()
    out[label] = {"whole": helpers.snapshot_whole(whole), "progress": hook.calls,
                  "type": type(whole).__name__}
print(json.dumps(out))
