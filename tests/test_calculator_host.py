"""Host logic of the drop-in calculator, checked against the compiled reference's outputs.

The device is replaced by tests/oracle_engine.OracleEngine so that this file runs without a
GPU; what is under test is everything *around* the kernel: feed rules (sort check, duplicate
rules, read-length sums), chromosome switching, result types and list lengths, read-less
chromosomes, missing tracks, skip_ncc, the -p N worker protocol (flush(chrom) + get_result).
Expected values come from the unmodified reference (tests/golden/ref_vectors.json).
The same scenarios run on the real device in tests/test_gpu_parity.py.
"""
import pickle

import numpy as np
import pytest

import helpers
from oracle_engine import OracleEngine
from pymasc_b200 import CCBitArrayCalculator, ReadUnsortedError, PositionOutOfRange, DeviceError


def feed_all(calc, toy, batch=False):
    for chrom, r in toy["reads"].items():
        if batch:
            calc.feed_batch(chrom, r["pos"], r["readlen"], r["is_reverse"])
            continue
        for p, l, rev in zip(r["pos"].tolist(), r["readlen"].tolist(), r["is_reverse"].tolist()):
            (calc.feed_reverse_read if rev else calc.feed_forward_read)(chrom, p, l)


def make_calc(spec, toy, engine=None):
    feeder = helpers.ListFeeder(toy["tracks"]) if toy["tracks"] is not None else None
    return CCBitArrayCalculator(spec["D"], spec["rl"], toy["references"], toy["lengths"], feeder,
                                spec["skip_ncc"], None, None, engine=engine or OracleEngine())


@pytest.mark.parametrize("batch", [False, True], ids=["per_read", "feed_batch"])
@pytest.mark.parametrize("spec", helpers.TOY_SPECS, ids=lambda s: s["name"])
def test_whole_result_equals_reference(spec, batch):
    toy = helpers.make_toy(spec)
    calc = make_calc(spec, toy)
    feed_all(calc, toy, batch)
    calc.finishup_calculation()
    got = helpers.snapshot_whole(calc.get_whole_result())
    helpers.assert_snapshots_equal(got, helpers.load_ref_vectors()["toy_" + spec["name"]])


@pytest.mark.parametrize("spec", helpers.TOY_SPECS[:3], ids=lambda s: s["name"])
def test_worker_protocol_equals_reference(spec):
    """handler/worker.py:224-234: flush(chrom) then get_result(chrom), one chromosome per
    calculator; results must pickle (they cross a multiprocessing.Queue)."""
    toy = helpers.make_toy(spec)
    want = helpers.load_ref_vectors()["toy_" + spec["name"] + "_worker"]
    for chrom in toy["references"]:
        calc = make_calc(spec, toy)
        if chrom in toy["reads"]:
            feed_all(calc, {"reads": {chrom: toy["reads"][chrom]}})
        calc.flush(chrom)
        res = pickle.loads(pickle.dumps(calc.get_result(chrom)))
        got = {"chrom": helpers.snapshot_chrom(res.chrom),
               "mappable_chrom": helpers.snapshot_chrom(res.mappable_chrom)}
        helpers.assert_snapshots_equal(got, want[chrom])


def test_fixture_through_calculator_equals_reference_and_golden():
    fx = helpers.load_fixture()
    vec = helpers.load_ref_vectors()
    for label, with_track, skip in (("fixture_ncc", False, False), ("fixture_both", True, False),
                                    ("fixture_mscc_only", True, True)):
        feeder = helpers.ListFeeder({"chr1": (fx["map_begins"], fx["map_ends"])}) if with_track else None
        calc = CCBitArrayCalculator(300, 36, fx["references"], fx["lengths"], feeder, skip,
                                    engine=OracleEngine())
        calc.feed_batch("chr1", fx["pos"], fx["readlen"], fx["is_reverse"])
        calc.finishup_calculation()
        got = helpers.snapshot_whole(calc.get_whole_result())
        want = vec[label]
        for k in ("chroms", "mappable_chroms"):
            if k in got:
                assert len(got[k]) == 86
                got[k] = {c: v for c, v in got[k].items() if c in want[k]}
        helpers.assert_snapshots_equal(got, want)
    gold = helpers.load_golden_tables()
    cc = np.array(got["mappable_chroms"]["chr1"]["cc"])
    assert np.array_equal(cc, np.array(gold["mscc"]["chr1"]))


def test_one_device_batch_for_the_whole_genome():
    """Closed chromosomes are queued and run as ONE batch together with the read-less
    chromosomes' autocorrelations."""
    spec = helpers.TOY_SPECS[0]
    toy = helpers.make_toy(spec)
    eng = OracleEngine()
    calc = make_calc(spec, toy, eng)
    feed_all(calc, toy)
    assert eng.batches == []
    calc.finishup_calculation()
    assert eng.batches == [3]          # c1, c3 with reads + c2's autocorrelation


def test_unsorted_position_raises():
    calc = CCBitArrayCalculator(50, 20, ["a", "b"], [1000, 1000], engine=OracleEngine())
    calc.feed_forward_read("a", 100, 20)
    with pytest.raises(ReadUnsortedError):
        calc.feed_reverse_read("a", 99, 20)
    with pytest.raises(ReadUnsortedError):
        calc.feed_batch("a", np.array([200, 150]), np.array([20, 20]), np.array([False, False]))


def test_returning_to_a_solved_chromosome_raises():
    calc = CCBitArrayCalculator(50, 20, ["a", "b"], [1000, 1000], engine=OracleEngine())
    calc.feed_forward_read("a", 100, 20)
    calc.feed_forward_read("b", 100, 20)
    with pytest.raises(ReadUnsortedError):    # 'a' is in _solved_chr (mscc.pyx:354-356)
        calc.feed_forward_read("a", 200, 20)


def test_unknown_chromosome_is_a_keyerror():
    calc = CCBitArrayCalculator(50, 20, ["a"], [1000], engine=OracleEngine())
    with pytest.raises(KeyError):
        calc.feed_forward_read("zzz", 1, 20)
    with pytest.raises(KeyError):
        calc.get_result("a")


def test_position_beyond_the_bit_array_raises_instead_of_aborting():
    calc = CCBitArrayCalculator(50, 20, ["a"], [1000], engine=OracleEngine())
    with pytest.raises(PositionOutOfRange):
        calc.feed_reverse_read("a", 1000 + 20 + 50 + 100, 20)
    feeder = helpers.ListFeeder({"a": (np.array([10]), np.array([5000]))})
    calc = CCBitArrayCalculator(50, 20, ["a"], [1000], feeder, engine=OracleEngine())
    calc.feed_forward_read("a", 5, 20)
    with pytest.raises(PositionOutOfRange):
        calc.finishup_calculation()


def test_readonly_attributes_and_constants():
    calc = CCBitArrayCalculator(50, 20, ["a"], [1000], None, True, "lock", engine=OracleEngine())
    assert (calc.max_shift, calc.read_len, calc.skip_ncc, calc.logger_lock) == (50, 20, True, "lock")
    assert calc.MAPPABILITY_THRESHOLD == 1.0 and calc.EXTRA_ALLOCATE_SIZE == 100
    with pytest.raises(AttributeError):
        calc.max_shift = 3


def test_default_engine_fails_loudly_without_a_gpu():
    """No CPU fallback: with no device (or no library) the first batch raises."""
    try:
        import torch
        if torch.cuda.is_available():
            pytest.skip("a GPU is present")
    except ImportError:
        pass
    calc = CCBitArrayCalculator(50, 20, ["a"], [1000])
    calc.feed_forward_read("a", 10, 20)
    calc.feed_reverse_read("a", 30, 20)
    with pytest.raises(DeviceError):
        calc.finishup_calculation()


def test_cc_from_the_count_vectors_is_bit_identical_to_calc_cc():
    """The calculator assigns result.cc from the int64 vectors (pymasc_b200.result.ncc_cc / mscc_cc)
    instead of calling calc_cc() on the list fields (PyMaSC/result.py:82-118): same float64 values,
    bit for bit -- NaN curves for all-zero counts, division by a zero mappable length, and the None
    slots of a mappable_len list when max_shift < 2 read_len - 1 included."""
    import numpy as np
    from pymasc_b200 import result as R
    rng = np.random.default_rng(5)
    for rl, D in ((36, 300), (36, 50), (36, 20), (5, 3), (101, 1300)):
        S = D + 1
        for zero in (False, True):
            ncc = np.zeros(S, np.int64) if zero else rng.integers(0, 5000, S).astype(np.int64)
            mscc = np.zeros(S, np.int64) if zero else rng.integers(0, 4000, S).astype(np.int64)
            mfwd = rng.integers(1, 90000, S).astype(np.int64)
            mrev = rng.integers(1, 90000, S).astype(np.int64)
            dmp = rng.integers(100000, 900000, S).astype(np.int64)
            dmp[rng.integers(0, S, 3)] = 0                        # division by zero somewhere
            L = 1000000
            n = R.NCCResult(max_shift=D, read_len=rl, genomelen=L, forward_sum=41234, reverse_sum=40111,
                            forward_read_len_sum=1, reverse_read_len_sum=1, ccbins=ncc.tolist())
            n.calc_cc()
            assert np.array_equal(R.ncc_cc(41234, 40111, ncc, L, D), n.cc, equal_nan=True)
            calc = CCBitArrayCalculator(D, rl, ["c"], [L], engine=OracleEngine())
            m = R.MSCCResult(max_shift=D, read_len=rl, genomelen=L, forward_sum=mfwd.tolist(), reverse_sum=mrev.tolist(),
                             forward_read_len_sum=1, reverse_read_len_sum=1, ccbins=mscc.tolist(),
                             mappable_len=calc._mappable_len_list(dmp))
            m.calc_cc()
            assert np.array_equal(R.mscc_cc(mfwd, mrev, mscc, dmp, rl, D), m.cc, equal_nan=True)
