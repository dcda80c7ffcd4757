"""Parity of the CUDA path (through the C ABI) against the CPU oracle and the reference's
vectors.  Integer counts must be bit-exact; the derived float64 curves are checked against
the reference's golden tables with the north_star tolerance (1e-9 relative, written below).
"""
import numpy as np
import pytest

import helpers
from oracle import cc_oracle
from pymasc_b200 import CCBitArrayCalculator, _shim, synth
from pymasc_b200.exceptions import PositionOutOfRange

pytestmark = pytest.mark.gpu
RTOL = 1e-9
PATHS = (1, 2, 3)   # 1 dense shift-window, 2 event-window, 3 lists (0 = auto picks one of them)
PATH_NAMES = {1: "dense", 2: "events", 3: "lists"}


@pytest.fixture(scope="module")
def dev():
    d = _shim.Device(0)
    d.set_option("h2d_min_group_bytes", 0)      # tests are small: let "h2d_groups" form groups anyway
    yield d
    d.close()


def check_job(dev, L, rl, D, fwd, rev, b=None, e=None, flags=None, lo=0, hi=0):
    """All three device algorithms -- the dense shift-window kernel (path 1), the event-window
    kernels (path 2) and the list kernels (path 3) -- against the oracle, bit for bit."""
    if flags is None:
        flags = _shim.FLAG_NCC | (_shim.FLAG_MSCC if b is not None else 0)
    mscc = bool(flags & _shim.FLAG_MSCC)
    want = cc_oracle.chrom_counts(L, rl, D, fwd, rev, b if mscc else None, e if mscc else None,
                                  skip_ncc=not flags & _shim.FLAG_NCC, lo=lo, hi=hi or None)
    try:
        for path in PATHS:
            if path == 3 and rl < 2:
                continue                    # the list path needs read_len >= 2 (auto never picks it then)
            dev.set_option("path", path)
            got = dev.run_batch([_shim.Job(L, rl, D, flags, fwd, rev, b, e, lo, hi)])[0]
            assert sorted(got) == sorted(want)
            for k in want:
                assert np.array_equal(np.asarray(got[k]), np.asarray(want[k])), (k, "path %d" % path)
    finally:
        dev.set_option("path", 0)
    return got


@pytest.fixture(params=PATHS, ids=lambda p: PATH_NAMES[p])
def pdev(dev, request):
    """The device pinned to one algorithm for the whole test."""
    dev.set_option("path", request.param)
    yield dev
    dev.set_option("path", 0)


def random_case(seed, L, rl, n_reads, with_track=True):
    rng = np.random.default_rng(seed)
    fwd = np.unique(rng.integers(1, L, size=n_reads)).astype(np.uint32)
    rev = np.unique(rng.integers(rl, L + rl - 1, size=n_reads)).astype(np.uint32)
    b = e = None
    if with_track:
        edges = np.sort(rng.choice(np.arange(0, L), size=2 * max(2, L // 700), replace=False))
        b, e = edges[0::2].astype(np.uint32), edges[1::2].astype(np.uint32)
    return fwd, rev, b, e


@pytest.mark.parametrize("L,rl,D", [
    (5000, 20, 100), (4097, 36, 127), (4096 - 36 - 128 - 100, 36, 128),   # word / unit edges
    (70001, 36, 300), (131072, 50, 30), (131072, 50, 80), (99999, 25, 49),
    (250000, 36, 1000), (60000, 101, 1300), (33, 5, 3),
])
def test_random_chromosomes_bit_exact(dev, L, rl, D):
    fwd, rev, b, e = random_case(L + D, L, rl, max(10, L // 20))
    check_job(dev, L, rl, D, fwd, rev, b, e)
    check_job(dev, L, rl, D, fwd, rev)                                   # NCC only
    check_job(dev, L, rl, D, fwd, rev, b, e, flags=_shim.FLAG_MSCC)      # skip_ncc


def test_all_kernel_variants_agree(dev):
    L, rl, D = 300000, 36, 500
    fwd, rev, b, e = random_case(5, L, rl, 40000)
    try:
        for v in range(6):
            dev.set_option("variant", v)
            check_job(dev, L, rl, D, fwd, rev, b, e)
            check_job(dev, L, rl, D, fwd, rev)
    finally:
        dev.set_option("variant", -1)


@pytest.mark.parametrize("tile", [64, 256, 2048])
def test_tile_sizes_agree(dev, tile):
    L, rl, D = 500000, 36, 700
    fwd, rev, b, e = random_case(6, L, rl, 30000)
    try:
        dev.set_option("tile_words", tile)
        check_job(dev, L, rl, D, fwd, rev, b, e)
    finally:
        dev.set_option("tile_words", 0)


def test_edge_inputs(dev):
    L, rl, D = 10000, 36, 200
    fwd, rev, b, e = random_case(9, L, rl, 500)
    empty = np.zeros(0, np.uint32)
    check_job(dev, L, rl, D, empty, rev, b, e)                 # no forward reads
    check_job(dev, L, rl, D, fwd, empty, b, e)                 # no reverse reads
    check_job(dev, L, rl, D, empty, empty, b, e)
    check_job(dev, L, rl, D, fwd, rev, empty, empty)           # track without intervals
    nbits = L + rl + D + 100
    # reads and intervals in the very last positions of the array; fully mappable chromosome
    check_job(dev, L, rl, D, np.array([1, L, nbits - 1], np.uint32),
              np.array([rl, L + rl - 1, nbits - 1], np.uint32),
              np.array([0], np.uint32), np.array([nbits - 1], np.uint32))
    # duplicates collapse
    check_job(dev, L, rl, D, np.array([5, 5, 5, 900], np.uint32), np.array([40, 40], np.uint32), b, e)
    # a long interval (interior handled by the long-fill kernel)
    check_job(dev, 400000, rl, D, fwd, rev, np.array([3, 300000], np.uint32),
              np.array([290000, 399000], np.uint32))


def test_out_of_range_is_an_error_code(dev):
    L, rl, D = 1000, 20, 50
    with pytest.raises(PositionOutOfRange):
        dev.run_batch([_shim.Job(L, rl, D, _shim.FLAG_NCC, [5], [L + rl + D + 100])])
    with pytest.raises(PositionOutOfRange):
        dev.run_batch([_shim.Job(L, rl, D, _shim.FLAG_NCC | _shim.FLAG_MSCC, [5], [30], [1], [5000])])


def test_position_shards_with_halos_add_up(dev):
    """Shards [lo, hi) with device-built halos sum to the whole (config 4's decomposition)."""
    L, rl, D = 600000, 36, 1000
    fwd, rev, b, e = random_case(11, L, rl, 60000)
    whole = check_job(dev, L, rl, D, fwd, rev, b, e)
    nbits = L + rl + D + 100
    cuts = [0, 150001, 300032, 449999, nbits]
    parts = [check_job(dev, L, rl, D, fwd, rev, b, e, lo=a, hi=z) for a, z in zip(cuts[:-1], cuts[1:])]
    for k in whole:
        assert np.array_equal(np.asarray(whole[k]), sum(np.asarray(p[k]) for p in parts)), k


def test_batch_of_mixed_jobs_and_autocorrelation(pdev):
    dev = pdev
    rl, D = 36, 300
    jobs, wants = [], []
    for i, L in enumerate([50000, 120000, 80000]):
        fwd, rev, b, e = random_case(20 + i, L, rl, 4000)
        if i == 1:
            jobs.append(_shim.Job(L, rl, D, _shim.FLAG_NCC, fwd, rev))
            wants.append(cc_oracle.chrom_counts(L, rl, D, fwd, rev))
        else:
            jobs.append(_shim.Job(L, rl, D, _shim.FLAG_NCC | _shim.FLAG_MSCC, fwd, rev, b, e))
            wants.append(cc_oracle.chrom_counts(L, rl, D, fwd, rev, b, e))
    _, _, b, e = random_case(30, 90000, rl, 10)
    jobs.append(_shim.Job(90000, rl, D, _shim.FLAG_AUTOCORR, None, None, b, e))
    wants.append({"dmp": cc_oracle.mappable_autocorr(90000, rl, D, b, e)})
    outs = dev.run_batch(jobs)
    for got, want in zip(outs, wants):
        assert sorted(got) == sorted(want)
        for k in want:
            assert np.array_equal(np.asarray(got[k]), np.asarray(want[k])), k


def test_large_max_shift_unit_chunks(dev):
    """max_shift 9000 -> 71 units -> two unit-chunk launches on moved base pointers."""
    L, rl, D = 200000, 36, 9000
    fwd, rev, b, e = random_case(40, L, rl, 20000)
    check_job(dev, L, rl, D, fwd, rev, b, e)


def test_prepare_execute_fetch_is_repeatable(dev):
    L, rl, D = 200000, 36, 400
    fwd, rev, b, e = random_case(41, L, rl, 20000)
    want = cc_oracle.chrom_counts(L, rl, D, fwd, rev, b, e)
    batch = dev.prepare([_shim.Job(L, rl, D, 3, fwd, rev, b, e)])
    for _ in range(3):
        batch.execute()
        got = batch.fetch()[0]
        for k in want:
            assert np.array_equal(np.asarray(got[k]), np.asarray(want[k])), k
    t = dev.timing()
    assert t["window_ms"] > 0 and t["launches"] >= 4
    batch.free()


def test_mappability_gather_and_its_fallback(dev):
    """Event path: M is gathered from sorted, disjoint intervals in one pass; anything else
    (overlaps, unsorted, end < begin) takes the scatter-style kernels behind the device-side
    check.  Both must give bit_array_set_region's result (bit_array.c:423-479)."""
    L, rl, D = 120000, 36, 300
    fwd, rev, b, e = random_case(80, L, rl, 8000)
    u32 = lambda x: np.asarray(x, np.uint32)
    # touching intervals, zero-length ones, a run of tiny ones (> 32 per 4096 bits), one huge
    tb = np.arange(20000, 24000, 10, dtype=np.uint32)
    sb = np.concatenate([u32([5, 100, 100, 300, 300]), tb, u32([30000, 119990])])
    se = np.concatenate([u32([100, 100, 300, 300, 1000]), tb + 4, u32([110000, L + rl + D + 99])])
    check_job(dev, L, rl, D, fwd, rev, sb, se)
    check_job(dev, L, rl, D, fwd, rev, sb[:1], se[:1])
    # overlapping, unsorted, and one with end < begin: the fallback
    ob = u32([5000, 100, 4000, 90000, 70000, 800])
    oe = u32([9000, 6000, 4500, 95000, 93000, 700])
    check_job(dev, L, rl, D, fwd, rev, ob, oe)
    check_job(dev, L, rl, D, fwd, rev, b[::-1].copy(), e[::-1].copy())
    # sharded ranges use the same build
    check_job(dev, L, rl, D, fwd, rev, sb, se, lo=30016, hi=90000)
    check_job(dev, L, rl, D, fwd, rev, ob, oe, lo=30016, hi=90000)
    # in one batch: a sorted job next to an unsorted one (the flag is per group, not per job)
    jobs = [_shim.Job(L, rl, D, 3, fwd, rev, sb, se), _shim.Job(L, rl, D, 3, fwd, rev, ob, oe),
            _shim.Job(L, rl, D, _shim.FLAG_AUTOCORR, None, None, ob, oe),
            _shim.Job(L, rl, D, _shim.FLAG_AUTOCORR, None, None, sb, se)]
    wants = [cc_oracle.chrom_counts(L, rl, D, fwd, rev, sb, se), cc_oracle.chrom_counts(L, rl, D, fwd, rev, ob, oe),
             {"dmp": cc_oracle.mappable_autocorr(L, rl, D, ob, oe)}, {"dmp": cc_oracle.mappable_autocorr(L, rl, D, sb, se)}]
    try:
        dev.set_option("path", 2)
        for groups in (1, 4):
            dev.set_option("h2d_groups", groups)
            for got, want in zip(dev.run_batch(jobs), wants):
                for k in want:
                    assert np.array_equal(np.asarray(got[k]), np.asarray(want[k])), (k, groups)
    finally:
        dev.set_option("path", 0)
        dev.set_option("h2d_groups", 4)


def test_packed_read_arrays_and_their_fallback(dev):
    """Event path: F and R are packed from the sorted reads by word-owning warps (no memset, no
    global atomics): long read-less runs (queued and zeroed by the whole grid), pile-ups that
    cross chunk boundaries, reads outside a shard's window -- and unsorted reads, which take the
    atomicOr scatter behind the device-side check.  bit_array_set_bit semantics (bit_array.c:668)."""
    rng = np.random.default_rng(123)
    L, rl, D = 3000000, 36, 300
    u32 = lambda x: np.asarray(x, np.uint32)
    # two clusters 2.5 Mb apart, a pile-up of 700 reads on 3 positions, a run of reads in one word
    left = np.sort(rng.integers(1, 200000, size=30000))
    right = np.sort(rng.integers(2800000, L, size=20000))
    pile = np.repeat(u32([150000, 150001, 150040]), [400, 250, 50])
    fwd = np.sort(np.concatenate([left, pile, right])).astype(np.uint32)
    rev = np.sort(np.concatenate([rng.integers(rl, 260000, size=25000), np.arange(2900000, 2900400),
                                  np.full(513, 2950000)])).astype(np.uint32)
    _, _, b, e = random_case(7, L, rl, 10)
    check_job(dev, L, rl, D, fwd, rev, b, e)
    check_job(dev, L, rl, D, fwd, rev, b, e, lo=100000 + 64, hi=2900000 - 32)     # reads outside the window
    check_job(dev, L, rl, D, fwd[:1], rev[-1:], b, e)
    check_job(dev, L, rl, D, fwd, rev)
    # unsorted reads: the fallback (the C ABI takes them; the calculator would have raised)
    perm = rng.permutation(fwd.size)
    check_job(dev, L, rl, D, fwd[perm], rev[::-1].copy(), b, e)
    # a batch that mixes sorted and unsorted jobs, with and without H2D groups
    jobs = [_shim.Job(L, rl, D, 3, fwd, rev, b, e), _shim.Job(L, rl, D, 3, fwd[perm], rev, b, e),
            _shim.Job(200000, rl, D, 1, left.astype(np.uint32), u32([]))]
    wants = [cc_oracle.chrom_counts(L, rl, D, fwd, rev, b, e)] * 2 + \
            [cc_oracle.chrom_counts(200000, rl, D, left.astype(np.uint32), u32([]))]
    try:
        dev.set_option("path", 2)
        for groups in (1, 3):
            dev.set_option("h2d_groups", groups)
            for got, want in zip(dev.run_batch(jobs), wants):
                for k in want:
                    assert np.array_equal(np.asarray(got[k]), np.asarray(want[k])), (k, groups)
    finally:
        dev.set_option("path", 0)
        dev.set_option("h2d_groups", 4)


@pytest.mark.parametrize("rl", [1, 32, 33, 129, 300, 1000, 1200])
def test_read_lengths_and_the_direct_lags(dev, rl):
    """The two directly counted lags (rl-1, rl-2) reach 1 .. 8 look-ahead quads of the gather;
    past that (rl 1200) the build falls back to fill + derive."""
    L, D = 90000, 200
    fwd, rev, b, e = random_case(90 + rl, L, rl, 3000)
    check_job(dev, L, rl, D, fwd, rev, b, e)


@pytest.mark.parametrize("groups", [1, 2, 3, 7])
def test_pipelined_input_copy_gives_the_same_counts(pdev, groups):
    """pmb_run_batch / async prepare: the H2D runs on a copy stream in groups of chromosomes and
    each group's build + kernels wait only for their own inputs.  Same counts as copy-then-compute,
    for every group count (more groups than jobs, jobs without reads or intervals included)."""
    dev = pdev
    rl, D = 36, 300
    jobs = []
    for i, L in enumerate([70000, 3000, 150000, 40000, 90000, 64]):
        fwd, rev, b, e = random_case(60 + i, L, rl, max(4, L // 12))
        if i == 1:
            fwd = fwd[:0]
        if i == 3:
            b, e = b[:0], e[:0]
        jobs.append(_shim.Job(L, rl, D, _shim.FLAG_NCC | _shim.FLAG_MSCC, fwd, rev, b, e))
    _, _, b, e = random_case(70, 50000, rl, 10)
    jobs.insert(2, _shim.Job(50000, rl, D, _shim.FLAG_AUTOCORR, None, None, b, e))
    dev.set_option("async_h2d", 0)
    sync = dev.prepare(jobs)
    sync.execute()
    want = sync.fetch()
    sync.free()
    k0 = jobs[0]
    ref = cc_oracle.chrom_counts(k0.chrom_len, rl, D, k0.fwd_pos, k0.rev_pos, k0.map_begins, k0.map_ends)
    for k in ref:
        assert np.array_equal(np.asarray(want[0][k]), np.asarray(ref[k])), k
    dev.set_option("h2d_groups", groups)
    try:
        dev.set_option("async_h2d", 1)
        batch = dev.prepare(jobs)
        for _ in range(2):                       # first execute is grouped, the second is not
            batch.execute()
            got = batch.fetch()
            for a, w in zip(got, want):
                assert sorted(a) == sorted(w)
                for k in w:
                    assert np.array_equal(np.asarray(a[k]), np.asarray(w[k])), k
        batch.free()
        dev.prepare(jobs).free()                 # never executed: free waits for the copy
        dev.set_option("async_h2d", 0)
        for a, w in zip(dev.run_batch(jobs), want):   # run_batch pipelines by itself
            for k in w:
                assert np.array_equal(np.asarray(a[k]), np.asarray(w[k])), k
    finally:
        dev.set_option("async_h2d", 0)
        dev.set_option("h2d_groups", 4)


def test_inputs_packed_in_one_host_buffer(pdev):
    """Position arrays that lie back to back in host memory with 16-byte padding travel as one
    copy per group; arrays that merely touch (no padding) or lie anywhere else do not merge."""
    dev = pdev
    rl, D = 36, 200
    cases = [random_case(200 + i, L, rl, max(5, L // 15)) for i, L in enumerate([30000, 30001, 50003, 777])]
    lens = [30000, 30001, 50003, 777]
    want = [cc_oracle.chrom_counts(L, rl, D, *cs) for L, cs in zip(lens, cases)]
    for pad in (True, False):
        total = sum(((a.size + 3) & ~3) if pad else a.size for cs in cases for a in cs)
        buf = np.zeros(total + 8, np.uint32)
        off, jobs = 0, []
        for L, cs in zip(lens, cases):
            views = []
            for a in cs:
                v = buf[off:off + a.size]
                v[:] = a
                views.append(v)
                off += ((a.size + 3) & ~3) if pad else a.size
            jobs.append(_shim.Job(L, rl, D, 3, *views))
        for groups in (1, 3):
            dev.set_option("h2d_groups", groups)
            try:
                got = dev.run_batch(jobs)
            finally:
                dev.set_option("h2d_groups", 4)
            for g, w in zip(got, want):
                for k in w:
                    assert np.array_equal(np.asarray(g[k]), np.asarray(w[k])), (k, pad, groups)


# ----------------------------------------------------------------- several read sets, one track
@pytest.mark.parametrize("groups", [1, 3])
def test_read_sets_sharing_one_track(pdev, groups):
    """BASELINE config 5 (multi-file run over one BigWig): FLAG_SAME_TRACK jobs read the M,
    Er/Or, edge sums and direct lags of the run's first job.  Every job must count exactly what
    it counts alone with its own copy of the intervals -- MSCC only, NCC+MSCC, with H2D groups,
    next to unrelated jobs, for sorted tracks (gather build + edge list) and for unsorted ones
    (scatter build + M scan fallback)."""
    dev = pdev
    rl, D = 36, 300
    S = _shim.FLAG_SAME_TRACK
    tracks, jobs, wants = [], [], []
    for c, (L, sorted_track) in enumerate([(90000, True), (150000, False), (40000, True)]):
        _, _, b, e = random_case(200 + c, L, rl, 10)
        if not sorted_track:
            b, e = b[::-1].copy(), e[::-1].copy()
        for f in range(3):
            fwd, rev, _, _ = random_case(210 + 10 * c + f, L, rl, max(8, L // (10 + 5 * f)), with_track=False)
            flags = _shim.FLAG_MSCC | (_shim.FLAG_NCC if (c + f) % 2 else 0)
            if f == 2 and c == 0:
                fwd = fwd[:0]                                    # a file without forward reads here
            jobs.append(_shim.Job(L, rl, D, flags | (S if f else 0), fwd, rev,
                                  None if f else b, None if f else e))
            wants.append(cc_oracle.chrom_counts(L, rl, D, fwd, rev, b, e, skip_ncc=not flags & _shim.FLAG_NCC))
        if c == 0:      # an unrelated NCC-only job and an autocorrelation between the runs
            fwd, rev, b2, e2 = random_case(300, 30000, rl, 900)
            jobs.append(_shim.Job(30000, rl, D, _shim.FLAG_NCC, fwd, rev))
            wants.append(cc_oracle.chrom_counts(30000, rl, D, fwd, rev))
            jobs.append(_shim.Job(30000, rl, D, _shim.FLAG_AUTOCORR, None, None, b2, e2))
            wants.append({"dmp": cc_oracle.mappable_autocorr(30000, rl, D, b2, e2)})
    dev.set_option("h2d_groups", groups)
    try:
        for got, want in zip(dev.run_batch(jobs), wants):
            assert sorted(got) == sorted(want)
            for k in want:
                assert np.array_equal(np.asarray(got[k]), np.asarray(want[k])), k
        # resident inputs, executed twice (the second execute is one ungrouped pass)
        dev.set_option("async_h2d", 1)
        batch = dev.prepare(jobs)
        for _ in range(2):
            batch.execute()
            for got, want in zip(batch.fetch(), wants):
                for k in want:
                    assert np.array_equal(np.asarray(got[k]), np.asarray(want[k])), k
        batch.free()
    finally:
        dev.set_option("async_h2d", 0)
        dev.set_option("h2d_groups", 4)


def test_shared_track_position_shards_and_short_reads(dev):
    """FLAG_SAME_TRACK with position shards (the runs share range_lo/range_hi) and with
    read_len 1 (no gathered M: the scan-based edge kernel runs, and must skip the followers)."""
    S = _shim.FLAG_SAME_TRACK
    for rl, D, lo, hi in ((36, 200, 20000 + 32, 61000), (1, 70, 0, 0)):
        L = 80000
        _, _, b, e = random_case(400 + rl, L, max(rl, 2), 10)
        jobs, wants = [], []
        for f in range(3):
            fwd, rev, _, _ = random_case(410 + f, L, max(rl, 2), 4000, with_track=False)
            jobs.append(_shim.Job(L, rl, D, 3 | (S if f else 0), fwd, rev, None if f else b, None if f else e, lo, hi))
            wants.append(cc_oracle.chrom_counts(L, rl, D, fwd, rev, b, e, lo=lo, hi=hi or None))
        try:
            for path in PATHS:
                if path == 3 and rl < 2:
                    continue
                dev.set_option("path", path)
                for got, want in zip(dev.run_batch(jobs), wants):
                    for k in want:
                        assert np.array_equal(np.asarray(got[k]), np.asarray(want[k])), (k, rl, path)
        finally:
            dev.set_option("path", 0)


def test_shared_track_flag_needs_a_matching_owner(dev):
    S = _shim.FLAG_SAME_TRACK
    fwd, rev, b, e = random_case(1, 5000, 20, 100)
    bad = [
        [_shim.Job(5000, 20, 50, 3 | S, fwd, rev)],                                            # first job
        [_shim.Job(5000, 20, 50, 1, fwd, rev), _shim.Job(5000, 20, 50, 3 | S, fwd, rev)],      # owner has no track
        [_shim.Job(5000, 20, 50, 3, fwd, rev, b, e), _shim.Job(6000, 20, 50, 3 | S, fwd, rev)],   # other chromosome
        [_shim.Job(5000, 20, 50, 3, fwd, rev, b, e), _shim.Job(5000, 20, 50, 3 | S, fwd, rev, None, None, 64, 2000)],
        [_shim.Job(5000, 20, 50, 3, fwd, rev, b, e), _shim.Job(5000, 20, 50, 1 | S, fwd, rev)],   # follower without MSCC
    ]
    for jobs in bad:
        with pytest.raises(ValueError):
            dev.run_batch(jobs)


def test_multifile_calculation_on_the_device():
    """MultiFileCalculation (one batch, shared tracks) against calculators finished alone."""
    from pymasc_b200 import MultiFileCalculation
    spec = helpers.TOY_SPECS[0]
    toys = [helpers.make_toy(dict(spec, seed=spec["seed"] + 100 * f)) for f in range(3)]
    feeder = helpers.ListFeeder(toys[0]["tracks"])
    del toys[2]["reads"]["c1"]
    alone, together = [], []
    for t in toys:
        for dst in (alone, together):
            c = CCBitArrayCalculator(spec["D"], spec["rl"], t["references"], t["lengths"], feeder, spec["skip_ncc"])
            for chrom, r in t["reads"].items():
                c.feed_batch(chrom, r["pos"], r["readlen"], r["is_reverse"])
            dst.append(c)
    for c in alone:
        c.finishup_calculation()
    MultiFileCalculation(together).finishup_calculation()
    for a, b in zip(alone, together):
        helpers.assert_snapshots_equal(helpers.snapshot_whole(b.get_whole_result()),
                                       helpers.snapshot_whole(a.get_whole_result()), rtol=RTOL)


# ----------------------------------------------------------------- the reference's own pins
def test_fixture_golden_tables_through_the_calculator():
    """configs[0]: tests/data ChIP-seq fixture, -d 300 -q 10 -r 36, against tests/golden."""
    fx = helpers.load_fixture()
    gold = helpers.load_golden_tables()
    vec = helpers.load_ref_vectors()
    feeder = helpers.ListFeeder({"chr1": (fx["map_begins"], fx["map_ends"])})
    calc = CCBitArrayCalculator(300, 36, fx["references"], fx["lengths"], feeder)
    calc.feed_batch("chr1", fx["pos"], fx["readlen"], fx["is_reverse"])
    calc.finishup_calculation()
    whole = calc.get_whole_result()
    assert (whole.forward_sum, whole.reverse_sum, whole.genomelen) == (622, 670, 3137454505)
    ncc, mscc = whole.chroms["chr1"], whole.mappable_chroms["chr1"]
    rows = gold["nreads"]["chr1"][1:]
    assert [int(x) for x in mscc.forward_sum] == [int(r.split("-")[0]) for r in rows]
    assert [int(x) for x in mscc.reverse_sum] == [int(r.split("-")[1]) for r in rows]
    assert mscc.mappable_len[35] == 19906
    for got, want in ((ncc.cc, gold["cc"]["chr1"]), (mscc.cc, gold["mscc"]["chr1"])):
        want = np.array(want)
        assert np.max(np.abs(np.asarray(got) - want) / np.abs(want)) <= RTOL
    got = helpers.snapshot_whole(whole)
    want = vec["fixture_both"]
    for k in ("chroms", "mappable_chroms"):
        got[k] = {c: v for c, v in got[k].items() if c in want[k]}
    helpers.assert_snapshots_equal(got, want, rtol=RTOL)


@pytest.mark.parametrize("spec", helpers.TOY_SPECS, ids=lambda s: s["name"])
def test_toy_cases_through_the_calculator(spec):
    toy = helpers.make_toy(spec)
    feeder = helpers.ListFeeder(toy["tracks"]) if toy["tracks"] is not None else None
    calc = CCBitArrayCalculator(spec["D"], spec["rl"], toy["references"], toy["lengths"], feeder,
                                spec["skip_ncc"])
    for chrom, r in toy["reads"].items():
        for p, l, rev in zip(r["pos"].tolist(), r["readlen"].tolist(), r["is_reverse"].tolist()):
            (calc.feed_reverse_read if rev else calc.feed_forward_read)(chrom, p, l)
    calc.finishup_calculation()
    helpers.assert_snapshots_equal(helpers.snapshot_whole(calc.get_whole_result()),
                                   helpers.load_ref_vectors()["toy_" + spec["name"]], rtol=RTOL)


# ----------------------------------------------------------------- full-size properties
def test_hg38_chr21_properties(pdev):
    dev = pdev
    """BASELINE config 2/3 shapes at full chromosome size (chr21 + chr22, max_shift 1000),
    through size-independent properties: dmp symmetry (popc(D_k) depends on |rl-1-k| only),
    k = 0 identities, shard additivity, and an oracle check on a window of shifts."""
    rl, D = 36, 1000
    # same read density as the 20 M-read genome
    share = (46709983 + 50818468) / 3088269832.0
    wl = synth.workload(n_reads=int(20_000_000 * share), chroms=["chr21", "chr22"], read_len=rl)
    jobs = [_shim.Job(w["length"], rl, D, 3, w["fwd"], w["rev"], w["begins"], w["ends"]) for w in wl]
    outs = dev.run_batch(jobs)
    for w, o in zip(wl, outs):
        dmp = o["dmp"]
        for k in range(0, 2 * rl - 1):
            assert dmp[k] == dmp[2 * (rl - 1) - k]                # |rl-1-k| symmetry
        assert o["fwd_total"] == np.unique(w["fwd"]).size
        assert o["rev_total"] == np.unique(w["rev"]).size
        assert np.all(o["mscc"] <= np.minimum(o["mfwd"], o["mrev"]))
        assert np.all(o["mscc"] <= o["ncc"]) and np.all(o["mfwd"] <= o["fwd_total"])
        # k = 0: ncc[0] = |F & R| computed independently on the host
        assert o["ncc"][0] == np.intersect1d(w["fwd"], w["rev"]).size
    # shard additivity at full size, 3 uneven shards of chr21
    w = wl[0]
    nbits = w["length"] + rl + D + 100
    cuts = [0, nbits // 3 + 5, 2 * nbits // 3 - 17, nbits]
    parts = dev.run_batch([_shim.Job(w["length"], rl, D, 3, w["fwd"], w["rev"], w["begins"], w["ends"], a, z)
                           for a, z in zip(cuts[:-1], cuts[1:])])
    for k in ("ncc", "dmp", "mfwd", "mrev", "mscc", "fwd_total", "rev_total"):
        assert np.array_equal(np.asarray(outs[0][k]), sum(np.asarray(p[k]) for p in parts)), k
    # oracle on the same full-size chromosome, restricted to a 2 Mb range so it takes seconds
    lo, hi = 10_000_000, 12_000_000
    got = dev.run_batch([_shim.Job(w["length"], rl, D, 3, w["fwd"], w["rev"], w["begins"], w["ends"], lo, hi)])[0]
    want = cc_oracle.chrom_counts(w["length"], rl, D, w["fwd"], w["rev"], w["begins"], w["ends"], lo=lo, hi=hi)
    for k in want:
        assert np.array_equal(np.asarray(got[k]), np.asarray(want[k])), k


# ----------------------------------------------------------------- the N > 1 decomposition
@pytest.mark.parametrize("world", [2, 8])
def test_rank_jobs_from_the_sharding_plan_add_up(pdev, world):
    dev = pdev
    """What bench.py / ShardedEngine run per rank: identical job lists with clipped inputs and
    empty ranges, executed here rank after rank on one device; the per-rank counter blocks
    must sum to the whole-genome result (the all-reduce is a plain int64 sum)."""
    from pymasc_b200 import sharding
    from pymasc_b200.distributed import shard_jobs
    rl, D = 36, 400
    chroms = []
    for i, L in enumerate([300000, 41000, 170000]):
        fwd, rev, b, e = random_case(50 + i, L, rl, L // 12)
        chroms.append({"length": L, "fwd": fwd, "rev": rev, "begins": b, "ends": e})
    lengths = [c["length"] for c in chroms]
    whole_jobs = sharding.jobs_for_rank(chroms, sharding.plan(lengths, rl, D, 1)[0], rl, D)
    _, _, b, e = random_case(60, 90000, rl, 10)
    whole_jobs.append(_shim.Job(90000, rl, D, _shim.FLAG_AUTOCORR, None, None, b, e))
    whole = dev.run_batch(whole_jobs)
    want_auto = cc_oracle.mappable_autocorr(90000, rl, D, b, e)
    assert np.array_equal(whole[-1]["dmp"], want_auto)
    parts = [dev.run_batch(shard_jobs(whole_jobs, r, world)) for r in range(world)]
    merged = sharding.sum_outputs(parts)
    for a, m in zip(whole, merged):
        assert sorted(a) == sorted(m)
        for k in a:
            assert np.array_equal(np.asarray(a[k]), np.asarray(m[k])), k


# ----------------------------------------------------------------- from the binary files
def test_files_to_golden_tables_on_the_device(tmp_path):
    """configs[0] end to end: the reference's BAM + BigWig fixtures through the native readers
    and the CUDA path (pymasc_b200.pipeline.run_files) against the reference's golden curves,
    and the mappable-length statistics (pymasc_b200.mappability) against the reference's JSON."""
    import json
    import os
    import shutil
    from pymasc_b200 import pipeline
    from pymasc_b200.mappability import MappabilityHandler
    gold = helpers.load_golden_tables()
    bam = os.path.join(helpers.GOLDEN, "ENCFF000RMB-test.bam")
    bw = str(tmp_path / "hg19_36mer-test.bigwig")
    shutil.copyfile(os.path.join(helpers.GOLDEN, "hg19_36mer-test.bigwig"), bw)
    results, info = pipeline.run_files([bam, bam], bw, max_shift=300, mapq_criteria=10)
    assert info["files"][0]["read_len"] == 36
    for whole in results:
        assert (whole.forward_sum, whole.reverse_sum, whole.genomelen) == (622, 670, 3137454505)
        for got, want in ((whole.chroms["chr1"].cc, gold["cc"]["chr1"]),
                          (whole.mappable_chroms["chr1"].cc, gold["mscc"]["chr1"])):
            want = np.array(want)
            assert np.max(np.abs(np.asarray(got) - want) / np.abs(want)) <= RTOL
    h = MappabilityHandler(bw, max_shift=300, readlen=36)
    h.calc_mappability()
    h.save_mappability_stats()
    assert json.load(open(h.map_path)) == gold["mappability_json"]


def test_multi_device_engine_with_two_contexts():
    """MultiDeviceEngine drives one pmb_ctx per device from its own host thread; with a single
    GPU in the box two contexts on device 0 exercise the same concurrency."""
    from pymasc_b200.multigpu import MultiDeviceEngine
    rl, D, S = 36, 300, _shim.FLAG_SAME_TRACK
    jobs, wants = [], []
    for c, L in enumerate([120000, 60000, 200000, 33000]):
        _, _, b, e = random_case(700 + c, L, rl, 10)
        for f in range(2):
            fwd, rev, _, _ = random_case(710 + 10 * c + f, L, rl, L // 15, with_track=False)
            jobs.append(_shim.Job(L, rl, D, 3 | (S if f else 0), fwd, rev, None if f else b, None if f else e))
            wants.append(cc_oracle.chrom_counts(L, rl, D, fwd, rev, b, e))
    eng = MultiDeviceEngine([0, 0])
    try:
        for _ in range(2):
            got = eng.run(jobs)
            assert all(len(idx) for idx in eng.last_placement)
            for g, w in zip(got, wants):
                for k in w:
                    assert np.array_equal(np.asarray(g[k]), np.asarray(w[k])), k
    finally:
        eng.close()
    assert _shim.device_count() >= 1


def test_multi_device_engine_position_split_with_two_contexts():
    from pymasc_b200.multigpu import MultiDeviceEngine
    rl, D = 36, 1300                       # two column chunks
    jobs, wants = [], []
    for c, L in enumerate([150000, 70000, 260000]):
        fwd, rev, b, e = random_case(800 + c, L, rl, L // 10)
        jobs.append(_shim.Job(L, rl, D, 3, fwd, rev, b, e))
        wants.append(cc_oracle.chrom_counts(L, rl, D, fwd, rev, b, e))
    eng = MultiDeviceEngine([0, 0, 0], split="position")
    try:
        got = eng.run(jobs)
    finally:
        eng.close()
    for g, w in zip(got, wants):
        for k in w:
            assert np.array_equal(np.asarray(g[k]), np.asarray(w[k])), k


# ----------------------------------------------------------------- full size against the oracle
def _density_workload(name, genome_reads, rl):
    total = float(sum(L for _, L in synth.HG38))
    L = dict(synth.HG38)[name]
    return synth.workload(n_reads=int(genome_reads * L / total), chroms=[name], read_len=rl)[0]


def test_whole_chr1_against_the_oracle_on_every_path_and_as_eight_shards(dev):
    """BASELINE configs[1]/[2] at full chromosome size: all of hg38 chr1 (249 Mbp, the 20 M-read
    genome's density, max_shift 1000) against cc_oracle -- every counter, every shift, on each of
    the three device algorithms -- and again as the sum of the 8 ranks' position shards
    (distributed.shard_jobs: what bench.py --gpus 8 and ShardedEngine run)."""
    from pymasc_b200 import sharding
    from pymasc_b200.distributed import shard_jobs
    rl, D = 36, 1000
    w = _density_workload("chr1", 20_000_000, rl)
    want = cc_oracle.chrom_counts(w["length"], rl, D, w["fwd"], w["rev"], w["begins"], w["ends"])
    job = _shim.Job(w["length"], rl, D, 3, w["fwd"], w["rev"], w["begins"], w["ends"])
    try:
        for path in PATHS:
            dev.set_option("path", path)
            got = dev.run_batch([job])[0]
            assert dev.timing()["path"] == path
            assert sorted(got) == sorted(want)
            for k in want:
                assert np.array_equal(np.asarray(got[k]), np.asarray(want[k])), (k, PATH_NAMES[path])
    finally:
        dev.set_option("path", 0)
    parts = [dev.run_batch(shard_jobs([job], r, 8)) for r in range(8)]
    merged = sharding.sum_outputs(parts)[0]
    for k in want:
        assert np.array_equal(np.asarray(merged[k]), np.asarray(want[k])), (k, "8 shards")


def test_deep_library_chromosome_at_max_shift_5000_against_the_oracle(dev):
    """BASELINE configs[3]'s shape: a 114 Mbp chromosome (chr13) at the 200 M-read library's
    density, max_shift 5000 -- five 1024-shift column chunks per event kernel, ~137 reverse reads
    inside every forward read's window on the list path -- against cc_oracle, paths 2 and 3, and
    the path `auto` picks."""
    rl, D = 36, 5000
    w = _density_workload("chr13", 200_000_000, rl)
    want = cc_oracle.chrom_counts(w["length"], rl, D, w["fwd"], w["rev"], w["begins"], w["ends"])
    job = _shim.Job(w["length"], rl, D, 3, w["fwd"], w["rev"], w["begins"], w["ends"])
    try:
        for path in (2, 3, 0):
            dev.set_option("path", path)
            got = dev.run_batch([job])[0]
            for k in want:
                assert np.array_equal(np.asarray(got[k]), np.asarray(want[k])), (k, path)
        assert dev.timing()["path"] == 3        # auto: ~28 pairs per read and 1024 shifts, under the limit of 64
    finally:
        dev.set_option("path", 0)


def test_auto_picks_the_documented_path_and_counts_stay_exact_across_the_thresholds(dev):
    """`path` 0 (what the product and bench.py run): lists while a forward read expects fewer
    than 64 reverse reads per 1024 shifts, else events unless the inputs hold more than 4 events
    per 32-bit word, else dense (include/pymasc_b200.h).  Inputs on both sides of each threshold;
    the path the library reports is asserted and the counts are checked against the oracle."""
    L, rl, D = 200000, 36, 1000
    words = (L + rl + D + 100) // 32
    rng = np.random.default_rng(77)

    def reads(n, lo, hi):
        return np.unique(rng.integers(lo, hi, size=int(n))).astype(np.uint32)

    _, _, b, e = random_case(5, L, rl, 10)
    cases = [   # (forward reads per word, reverse reads per word, read_len, expected path)
        (0.05, 0.05, rl, 3),        # ChIP-seq density
        (0.3, 1.8, rl, 3),          # ~56 pairs per read: still lists
        (0.3, 2.4, rl, 2),          # ~75 pairs per read, 2.7 events per word: events
        (0.3, 3.4, rl, 2),          # 3.7 events per word + run edges: just under 4
        (2.2, 2.6, rl, 1),          # 4.8 events per word: dense
        (0.05, 0.05, 1, 2),         # read_len 1: no list path
    ]
    try:
        for f_pw, r_pw, rlen, expect in cases:
            fwd = reads(f_pw * words * 1.02, 1, L)
            rev = reads(r_pw * words * 1.02, rlen, L + rlen - 1)
            dev.set_option("path", 0)
            got = dev.run_batch([_shim.Job(L, rlen, D, 3, fwd, rev, b, e)])[0]
            assert dev.timing()["path"] == expect, (f_pw, r_pw, rlen, fwd.size, rev.size)
            want = cc_oracle.chrom_counts(L, rlen, D, fwd, rev, b, e)
            for k in want:
                assert np.array_equal(np.asarray(got[k]), np.asarray(want[k])), (k, f_pw, r_pw)
    finally:
        dev.set_option("path", 0)


@pytest.mark.parametrize("cap", [1, 37, 1000])
def test_list_pair_histogram_is_flushed_before_its_16_bit_halves_can_overflow(dev, cap):
    """List path, NCC+MSCC: ncc and mscc of a shift share one 32-bit shared-memory bin (low / high
    half), a read adds at most 1 to a bin, and a CTA's segment is cut after 65535 reads
    (pmb_lists.cuh, LIST_PAIR_CAP).  The cut is exercised here with a tiny cap (option
    "list_pair_cap"): same counts as the oracle whatever the cap, pile-ups on one shift
    included (every forward read has its partner at distance 150)."""
    L, rl, D = 400000, 36, 300
    fwd, rev, b, e = random_case(4242 + cap, L, rl, 60000)
    rev = np.unique(np.concatenate([rev, fwd[fwd + 150 < L + rl - 1] + 150])).astype(np.uint32)
    want = cc_oracle.chrom_counts(L, rl, D, fwd, rev, b, e)
    assert int(want["ncc"][150]) >= fwd.size - 200          # the pile-up is there
    try:
        dev.set_option("path", 3)
        dev.set_option("list_pair_cap", cap)
        jobs = [_shim.Job(L, rl, D, 3, fwd, rev, b, e), _shim.Job(L, rl, D, 2, fwd, rev, b, e),
                _shim.Job(L, rl, D, 1, fwd, rev)]
        outs = dev.run_batch(jobs)
        assert dev.timing()["path"] == 3
        for k in want:
            assert np.array_equal(np.asarray(outs[0][k]), np.asarray(want[k])), k
        for k in ("mscc", "mfwd", "mrev", "dmp"):
            assert np.array_equal(np.asarray(outs[1][k]), np.asarray(want[k])), k
        assert np.array_equal(np.asarray(outs[2]["ncc"]), np.asarray(want["ncc"]))
    finally:
        dev.set_option("list_pair_cap", 0)
        dev.set_option("path", 0)


def test_unsorted_input_moves_a_list_batch_to_the_event_path(dev):
    """The list path verifies sortedness on the device while it counts; unsorted reads or
    overlapping intervals send the batch to the event path (fetch, or pmb_batch_validate for
    callers that read the device counters without a fetch) -- same counts, path 2 reported."""
    L, rl, D = 150000, 36, 300
    fwd, rev, b, e = random_case(31, L, rl, 9000)
    want = cc_oracle.chrom_counts(L, rl, D, fwd, rev, b, e)
    perm = np.random.default_rng(1).permutation(fwd.size)
    ob, oe = np.asarray([5000, 100, 4000], np.uint32), np.asarray([9000, 6000, 4500], np.uint32)
    want_o = cc_oracle.chrom_counts(L, rl, D, fwd, rev, ob, oe)
    try:
        dev.set_option("path", 3)
        got = dev.run_batch([_shim.Job(L, rl, D, 3, fwd, rev, b, e)])[0]
        assert dev.timing()["path"] == 3
        for jobs, w in (([_shim.Job(L, rl, D, 3, fwd[perm], rev, b, e)], want),
                        ([_shim.Job(L, rl, D, 3, fwd, rev, ob, oe)], want_o)):
            got = dev.run_batch(jobs)[0]
            assert dev.timing()["path"] == 2
            for k in w:
                assert np.array_equal(np.asarray(got[k]), np.asarray(w[k])), k
            batch = dev.prepare(jobs)           # the all-reduce route: execute, validate, then read
            batch.execute()
            assert dev.timing()["path"] == 3
            batch.validate()
            assert dev.timing()["path"] == 2
            batch.execute()                     # later executes stay on the event path
            assert dev.timing()["path"] == 2
            got = batch.fetch()[0]
            batch.free()
            for k in w:
                assert np.array_equal(np.asarray(got[k]), np.asarray(w[k])), k
    finally:
        dev.set_option("path", 0)


@pytest.mark.parametrize("groups", [1, 3])
def test_repeated_executes_are_bit_identical(pdev, groups):
    """Every count is an exact integer sum, so whatever order the atomics land in, two executes
    of the same resident batch must agree bit for bit: a data race between warps, streams (the
    grouped first execute runs on two) or kernels, or a read of memory an earlier execute left
    behind, shows up as a difference (compute-sanitizer's racecheck / initcheck are closed on this
    pool).  30 executes of a mixed batch per path, the first one grouped."""
    dev = pdev
    rl, D, S = 36, 700, _shim.FLAG_SAME_TRACK
    jobs = []
    for c, L in enumerate([180000, 52000, 240000, 9000]):
        fwd, rev, b, e = random_case(900 + c, L, rl, L // 9)
        jobs.append(_shim.Job(L, rl, D, 3 if c != 1 else 1, fwd, rev, b if c != 1 else None, e if c != 1 else None))
        if c == 2:
            fwd2, rev2, _, _ = random_case(950, L, rl, L // 14, with_track=False)
            jobs.append(_shim.Job(L, rl, D, 3 | S, fwd2, rev2))
    _, _, b, e = random_case(960, 70000, rl, 10)
    jobs.append(_shim.Job(70000, rl, D, _shim.FLAG_AUTOCORR, None, None, b, e))
    dev.set_option("h2d_groups", groups)
    dev.set_option("async_h2d", 1)
    try:
        batch = dev.prepare(jobs)
        first = None
        for it in range(30):
            batch.execute()
            got = batch.fetch()
            if first is None:
                first = got
                continue
            for a, w in zip(got, first):
                for k in w:
                    assert np.array_equal(np.asarray(a[k]), np.asarray(w[k])), (k, it)
        batch.free()
    finally:
        dev.set_option("async_h2d", 0)
        dev.set_option("h2d_groups", 4)
