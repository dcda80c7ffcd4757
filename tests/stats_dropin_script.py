"""Run in a fresh interpreter by tests/test_stats_dropin.py: the B200 calculator INSIDE the
reference package (oracle/_ref/PyMaSC importable, so pymasc_b200.result hands back the
reference's own result classes), followed by the reference's unmodified stats.py.  Prints the
genome-wide statistics as JSON.  argv[1]: "oracle" (CPU checker engine) or "device"."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "oracle", "ref_stubs"), os.path.join(ROOT, "oracle", "_ref"),
          os.path.join(ROOT, "tests"), ROOT):
    sys.path.insert(0, p)

import PyMaSC.stats as ref_stats                      # noqa: E402  the reference, unmodified
import PyMaSC.result as ref_result                    # noqa: E402
import helpers                                        # noqa: E402
from pymasc_b200 import CCBitArrayCalculator, io_native, result as our_result   # noqa: E402

assert our_result.USING_REFERENCE_TYPES

engine = None
if sys.argv[1] == "oracle":
    from oracle_engine import OracleEngine
    engine = OracleEngine()


class Config(object):                                 # the CLI defaults (utils/parsearg.py:241-258)
    read_length = 36
    chi2_pval = 0.05
    mv_avr_filter_len = 15
    filter_mask_len = 5
    min_calc_width = 50
    expected_library_length = None


bam = io_native.BamFile(os.path.join(helpers.GOLDEN, "ENCFF000RMB-test.bam"), mapq_criteria=10)
bw = io_native.BigWigReader(os.path.join(helpers.GOLDEN, "hg19_36mer-test.bigwig"))
calc = CCBitArrayCalculator(300, 36, list(bam.references), list(bam.lengths), bw, engine=engine)
io_native.feed_bam(calc, bam)
calc.finishup_calculation()
whole = calc.get_whole_result()
assert isinstance(whole, ref_result.BothGenomeWideResult)
stats = ref_stats.make_genome_wide_stat(whole, Config(), False)
out = {}
for key, w in (("ncc", stats.whole_ncc_stats), ("mscc", stats.whole_mscc_stats)):
    m = w.stats.metrics_at_estimated_length
    out[key] = {"est_lib_len": int(w.est_lib_len), "fragment_length": int(m.fragment_length),
                "ccfl": float(m.ccfl), "fwhm": int(m.fwhm), "nsc": float(m.nsc), "rsc": float(m.rsc),
                "vsn": float(m.vsn), "cc_min": float(w.stats.cc_min), "ccrl": float(w.stats.ccrl)}
out["forward_reads"] = int(stats.whole_ncc_stats.stats.forward_reads)
out["reverse_reads"] = int(stats.whole_ncc_stats.stats.reverse_reads)
out["chrom_types"] = sorted({type(v).__name__ for v in whole.chroms.values()})
print(json.dumps(out))
