"""The list path's arithmetic (oracle/list_model.py, the thread-for-thread restatement of
pymasc_b200/csrc/pmb_lists.cuh) against the pinned C oracle -- runs without a GPU."""
import numpy as np
import pytest

from oracle import cc_oracle, list_model


def random_case(seed, L, rl, n_reads, n_runs=None):
    rng = np.random.default_rng(seed)
    fwd = np.sort(rng.integers(1, L, size=n_reads)).astype(np.uint32)          # duplicates stay in
    rev = np.sort(rng.integers(rl, L + rl - 1, size=n_reads)).astype(np.uint32)
    n_runs = n_runs or max(2, L // 300)
    edges = np.sort(rng.choice(np.arange(0, L), size=2 * n_runs, replace=False))
    return fwd, rev, edges[0::2].astype(np.uint32), edges[1::2].astype(np.uint32)


def assert_same(got, want):
    assert sorted(got) == sorted(want)
    for k in want:
        assert np.array_equal(np.asarray(got[k]), np.asarray(want[k])), k


@pytest.mark.parametrize("L,rl,D,n", [
    (3000, 20, 100, 400), (5000, 36, 127, 900), (2000, 5, 3, 300), (9000, 36, 300, 700),
    (4000, 50, 30, 2500), (6000, 25, 49, 100), (2500, 2, 64, 500), (700, 101, 1300, 200),
])
def test_list_model_equals_the_oracle(L, rl, D, n):
    fwd, rev, b, e = random_case(L + D, L, rl, n)
    assert_same(list_model.chrom_counts(L, rl, D, fwd, rev, b, e), cc_oracle.chrom_counts(L, rl, D, fwd, rev, b, e))
    assert_same(list_model.chrom_counts(L, rl, D, fwd, rev), cc_oracle.chrom_counts(L, rl, D, fwd, rev))
    assert_same(list_model.chrom_counts(L, rl, D, fwd, rev, b, e, skip_ncc=True),
                cc_oracle.chrom_counts(L, rl, D, fwd, rev, b, e, skip_ncc=True))


def test_list_model_position_shards():
    L, rl, D = 8000, 36, 250
    fwd, rev, b, e = random_case(3, L, rl, 1500)
    nbits = L + rl + D + 100
    for lo, hi in ((0, 2001), (2001, 4032), (4032, 7999), (7999, nbits), (100, 101), (5000, 5000)):
        assert_same(list_model.chrom_counts(L, rl, D, fwd, rev, b, e, lo=lo, hi=hi),
                    cc_oracle.chrom_counts(L, rl, D, fwd, rev, b, e, lo=lo, hi=hi))


def test_list_model_interval_shapes():
    L, rl, D = 6000, 36, 200
    fwd, rev, _, _ = random_case(4, L, rl, 900)
    u32 = lambda x: np.asarray(x, np.uint32)
    nbits = L + rl + D + 100
    tb = np.arange(2000, 2400, 10, dtype=np.uint32)
    cases = [
        (np.concatenate([u32([5, 100, 100, 300, 300]), tb, u32([3000, 5990])]),       # touching, empty, tiny
         np.concatenate([u32([100, 100, 300, 300, 1000]), tb + 4, u32([5500, nbits - 1])])),
        (u32([0]), u32([nbits - 1])),                                                 # everything mappable
        (u32([]), u32([])),
        (u32([10]), u32([11])),
        (np.arange(0, 5000, 2, dtype=np.uint32), np.arange(1, 5001, 2, dtype=np.uint32)),   # every other bit
    ]
    for b, e in cases:
        assert_same(list_model.chrom_counts(L, rl, D, fwd, rev, b, e), cc_oracle.chrom_counts(L, rl, D, fwd, rev, b, e))
        assert_same(list_model.chrom_counts(L, rl, D, fwd, rev, b, e, lo=1000, hi=3000),
                    cc_oracle.chrom_counts(L, rl, D, fwd, rev, b, e, lo=1000, hi=3000))
        assert np.array_equal(list_model.mappable_autocorr(L, rl, D, b, e), cc_oracle.mappable_autocorr(L, rl, D, b, e))


def test_list_model_last_positions_and_duplicates():
    L, rl, D = 1000, 36, 200
    nbits = L + rl + D + 100
    u32 = lambda x: np.asarray(x, np.uint32)
    b, e = u32([0]), u32([nbits - 1])
    fwd, rev = u32([1, L, nbits - 1]), u32([rl, L + rl - 1, nbits - 1])
    assert_same(list_model.chrom_counts(L, rl, D, fwd, rev, b, e), cc_oracle.chrom_counts(L, rl, D, fwd, rev, b, e))
    fwd, rev = u32([5, 5, 5, 900]), u32([40, 40])
    assert_same(list_model.chrom_counts(L, rl, D, fwd, rev, b, e), cc_oracle.chrom_counts(L, rl, D, fwd, rev, b, e))
