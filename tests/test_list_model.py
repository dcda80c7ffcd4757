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


# ---- the devices the kernels use where they carry risk (oracle/list_model.py's header) ----
@pytest.mark.parametrize("cap", [1, 7, 37, 1000])
def test_packed_pair_histogram_with_any_segment_cap(cap):
    """ncc and mscc of a shift share one 32-bit bin; a read adds at most 1 to a bin and the bins
    are flushed every `cap` reads (LIST_PAIR_CAP = 65535 on the device): same counts whatever
    the cap, duplicates and a pile-up on one shift included."""
    L, rl, D = 5000, 36, 200
    fwd, rev, b, e = random_case(77 + cap, L, rl, 1200)
    rev = np.sort(np.concatenate([rev, fwd[fwd + 150 < L + rl - 1] + 150])).astype(np.uint32)
    want = cc_oracle.chrom_counts(L, rl, D, fwd, rev, b, e)
    assert_same(list_model.chrom_counts(L, rl, D, fwd, rev, b, e, pair_cap=cap), want)
    assert int(want["ncc"][150]) > 900


def test_packed_pair_histogram_needs_its_cap(monkeypatch):
    """The negative control: 66 000 reads with a partner at the same distance and no flush in
    between carry the low half of that bin into the high half -- the model keeps the bins modulo
    2^32 as the device does, so it shows; with the cap in place the counts are right."""
    L, rl, D = 300000, 36, 40
    fwd = np.arange(10, 10 + 4 * 66000, 4, dtype=np.uint32)
    rev = (fwd + 5).astype(np.uint32)
    b, e = np.asarray([0], np.uint32), np.asarray([L + 50], np.uint32)
    want = cc_oracle.chrom_counts(L, rl, D, fwd, rev, b, e)
    assert int(want["ncc"][5]) == 66000
    assert_same(list_model.chrom_counts(L, rl, D, fwd, rev, b, e), want)
    monkeypatch.setattr(list_model, "PAIR_CAP", 1 << 30)
    got = list_model.chrom_counts(L, rl, D, fwd, rev, b, e, pair_cap=1 << 30)
    assert int(got["ncc"][5]) != 66000 or int(got["mscc"][5]) != int(want["mscc"][5])


def test_cursor_searches_equal_bisect():
    """cursor_window + the group's vote + gallop / binary search (pmb_lists.cuh: cursor_window,
    cursor_resolve) against bisect, on lists shorter than the window, clustered lists, long gaps
    and duplicates, with the cursor anywhere at or before the first key's position."""
    import bisect
    rng = np.random.default_rng(11)
    for trial in range(300):
        n = int(rng.choice([0, 1, 5, 62, 63, 64, 200, 3000]))
        kind = trial % 3
        if kind == 0:
            a = np.sort(rng.integers(0, max(1, 4 * n), size=n))
        elif kind == 1:
            a = np.sort(np.concatenate([rng.integers(0, 50, size=n // 2), rng.integers(10 ** 6, 10 ** 6 + 50, size=n - n // 2)]))
        else:
            a = np.sort(rng.integers(0, 10, size=n) * 1000)
        a = [int(x) for x in a]
        top = (a[-1] if a else 0) + 100
        keys = sorted(int(x) for x in rng.integers(0, top, size=int(rng.integers(1, 33))))
        first = bisect.bisect_left(a, keys[0])
        cur = int(rng.integers(0, first + 1)) if trial % 4 else first
        # precondition of the kernels' cursor: everything before it is below every key of the group
        cur = min(cur, first)
        got = list_model.group_lower_bounds(a, len(a), cur, keys, keys[-1])
        assert got == [bisect.bisect_left(a, k) for k in keys], (trial, n, cur)


def test_run_edge_distances_modulo_2_32():
    """The run-edge kernel takes distances modulo 2^32 and compares them with max_shift as unsigned
    numbers.  Exact for every pair of positions of a chromosome that keeps 2 (max_shift + read_len)
    + 64 bits clear of 2^32 (what pmb_api.cu requires of the list path) -- and NOT beyond: the
    second half is the counter-example the margin exists for."""
    rng = np.random.default_rng(5)
    for D, rl in ((1000, 36), (5000, 101), (17000, 2), (200, 300)):
        nbits = (1 << 32) - 2 * (D + rl) - 64 - 1
        for off in (rl - 1 - D, 0):                      # cross-correlation windows, autocorrelation
            pos = np.concatenate([rng.integers(1, nbits, size=400), [1, 2, nbits - 1, nbits, D, D + rl]])
            for own in (int(x) for x in pos[:60]):       # s or t of the thread's own run
                base = list_model.u32(own + off)
                for edge in (int(x) for x in pos):
                    exact = edge - (own + off)
                    want = exact if 0 <= exact <= D else None
                    assert list_model.edge_in_window(list_model.u32(edge), base, D) == want
    # beyond the margin a huge true distance wraps into the window
    D, rl = 1000, 36
    own, edge = 1, (1 << 32) - 2
    exact = edge - (own + rl - 1 - D)
    assert exact > D and list_model.edge_in_window(edge, list_model.u32(own + rl - 1 - D), D) is not None


def test_list_model_on_many_small_random_cases():
    """Short lists (every search falls back to gallop + binary search), no reads on a strand, no
    intervals, max_shift 0 and beyond the chromosome, read_len 2 ... beyond max_shift, shards."""
    rng = np.random.default_rng(2026)
    for trial in range(60):
        L = int(rng.choice([40, 150, 700, 2500]))
        rl = int(rng.choice([2, 3, 36, 75, 160]))
        D = int(rng.choice([0, 1, 35, 71, 149, 400]))
        n = int(rng.choice([0, 1, 3, 40, 300]))
        fwd = np.sort(rng.integers(1, L, size=n)).astype(np.uint32)
        rev = np.sort(rng.integers(rl, L + rl - 1, size=int(rng.choice([0, 1, n])))).astype(np.uint32)
        n_runs = int(rng.choice([0, 1, 2, max(2, L // 30)]))
        edges = np.sort(rng.choice(np.arange(0, L), size=2 * n_runs, replace=False))
        b, e = edges[0::2].astype(np.uint32), edges[1::2].astype(np.uint32)
        nbits = L + rl + D + 100
        lo = int(rng.integers(0, nbits)) if trial % 3 == 0 else 0
        hi = int(rng.integers(lo, nbits + 1)) if trial % 3 == 0 else None
        want = cc_oracle.chrom_counts(L, rl, D, fwd, rev, b, e, lo=lo, hi=hi)
        assert_same(list_model.chrom_counts(L, rl, D, fwd, rev, b, e, lo=lo, hi=hi, pair_cap=int(rng.choice([1, 5, 65535]))), want)
        assert np.array_equal(list_model.mappable_autocorr(L, rl, D, b, e), cc_oracle.mappable_autocorr(L, rl, D, b, e))
