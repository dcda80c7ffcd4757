"""pmb_run_chrom_noalloc (include/pymasc_b200.h): the allocation-free entry point -- scratch in a
static __device__ array of the library -- against the existing entry point pmb_run_batch, through
the C ABI.  Integer counts: bit-exact (the tolerance of every count test here is ==).

Collected last on purpose (the name): a failure here hides
nothing behind it.  Passes alone in a fresh process and after the other GPU tests: the scratch
array keeps whatever the previous call (of any test) left there, and every call clears it.
"""
import numpy as np
import pytest

from pymasc_b200 import _shim

pytestmark = pytest.mark.gpu


def _case(seed, L, rl, n_reads):
    rng = np.random.default_rng(seed)
    fwd = np.sort(rng.integers(1, L, size=n_reads)).astype(np.uint32)            # duplicates stay in
    rev = np.sort(rng.integers(rl, L + rl - 1, size=n_reads)).astype(np.uint32)
    edges = np.sort(rng.choice(np.arange(0, L), size=2 * max(2, L // 700), replace=False))
    return fwd, rev, edges[0::2].astype(np.uint32), edges[1::2].astype(np.uint32)


def _noalloc(dev, torch, L, rl, D, flags, fwd, rev, b, e, lo=0, hi=0):
    """One call of the new entry point: the caller (this test) owns inputs and outputs in device
    memory; the entry point itself allocates nothing."""
    def up(a):
        a = np.ascontiguousarray(a if a is not None else np.zeros(0, np.uint32), dtype=np.uint32)
        t = torch.from_numpy(a.view(np.int32).copy()).cuda() if a.size else None
        return t, (t.data_ptr() if t is not None else 0), a.size
    tf, pf, nf = up(fwd)
    tr, pr, nr = up(rev)
    tb, pb, nb = up(b)
    te, pe, _ = up(e)
    n_out = dev.noalloc_out_int64(D)
    out = torch.full((n_out,), -7, dtype=torch.int64, device="cuda")      # garbage: the call must clear it
    torch.cuda.synchronize()
    dev.run_chrom_noalloc(L, rl, D, flags, pf, nf, pr, nr, pb, pe, nb, out.data_ptr(), n_out, lo, hi)
    dev.synchronize()
    h = out.cpu().numpy()
    S = D + 1
    res = {name: h[i * S:(i + 1) * S] for i, name in enumerate(_shim.COUNTER_ORDER)}
    res["fwd_total"], res["rev_total"] = int(h[5 * S]), int(h[5 * S + 1])
    status = h[5 * S + 2:5 * S + 3].view(np.int32)
    del tf, tr, tb, te
    return res, (int(status[0]), int(status[1]))


def _same(got, want):
    for k, v in want.items():
        assert np.array_equal(np.asarray(got[k]), np.asarray(v)), k


def test_noalloc_entry_point_equals_the_existing_one_twice_in_a_row():
    torch = pytest.importorskip("torch")
    dev = _shim.Device(0)
    try:
        rl = 36
        # two calls in a row on different inputs (the second must not see the first one's sums),
        # then the first again; NCC+MSCC, MSCC only on a position shard, a larger max_shift
        calls = [
            (250000, 1000, 3, _case(1, 250000, rl, 30000), 0, 0),
            (90000, 300, _shim.FLAG_MSCC, _case(2, 90000, rl, 9000), 20000 + 32, 61000),
            (250000, 1000, 3, _case(1, 250000, rl, 30000), 0, 0),
            (120000, 4095, 3, _case(3, 120000, rl, 15000), 0, 0),
            (60000, 200, _shim.FLAG_NCC, _case(4, 60000, rl, 5000), 0, 0),
        ]
        for L, D, flags, (fwd, rev, b, e), lo, hi in calls:
            mscc = bool(flags & _shim.FLAG_MSCC)
            want = dev.run_batch([_shim.Job(L, rl, D, flags, fwd, rev, b if mscc else None,
                                            e if mscc else None, lo, hi)])[0]
            got, status = _noalloc(dev, torch, L, rl, D, flags, fwd, rev, b if mscc else None,
                                   e if mscc else None, lo, hi)
            assert status == (0, 0)
            _same(got, want)
            for k in _shim.COUNTER_ORDER:            # vectors the flags do not select are zero
                if k not in want:
                    assert not got[k].any(), k
        # read-less chromosome: the mappable-length autocorrelation (mscc.pyx:207-215)
        _, _, b, e = _case(5, 80000, rl, 10)
        want = dev.run_batch([_shim.Job(80000, rl, 400, _shim.FLAG_AUTOCORR, None, None, b, e)])[0]
        got, status = _noalloc(dev, torch, 80000, rl, 400, _shim.FLAG_AUTOCORR, None, None, b, e)
        assert status == (0, 0)
        assert np.array_equal(got["dmp"], want["dmp"])
    finally:
        dev.close()


def test_noalloc_entry_point_refuses_what_its_scratch_cannot_hold():
    torch = pytest.importorskip("torch")
    dev = _shim.Device(0)
    try:
        out = torch.zeros(dev.noalloc_out_int64(5000), dtype=torch.int64, device="cuda")
        with pytest.raises(ValueError, match="PMB_NOALLOC_MAX_SHIFT"):
            dev.run_chrom_noalloc(10000, 36, 4096, 1, 0, 0, 0, 0, 0, 0, 0, out.data_ptr(), out.numel())
        with pytest.raises(ValueError, match="int64"):                       # output block too small
            dev.run_chrom_noalloc(10000, 36, 1000, 1, 0, 0, 0, 0, 0, 0, 0, out.data_ptr(), 100)
        with pytest.raises(ValueError):                                      # PMB_FLAG_SAME_TRACK
            dev.run_chrom_noalloc(10000, 36, 100, 3 | _shim.FLAG_SAME_TRACK, 0, 0, 0, 0, 0, 0, 0,
                                  out.data_ptr(), out.numel())
        # unsorted reads are reported in the status word, not silently miscounted
        fwd = np.array([500, 40, 900], np.uint32)
        got, status = _noalloc(dev, torch, 10000, 36, 100, 1, fwd, np.array([700], np.uint32), None, None)
        assert status[0] == 0 and status[1] != 0
        # a position beyond the array: the range flag
        got, status = _noalloc(dev, torch, 1000, 36, 100, 1, np.array([5], np.uint32),
                               np.array([1000 + 36 + 100 + 100], np.uint32), None, None)
        assert status[0] != 0
    finally:
        dev.close()
