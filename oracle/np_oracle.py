"""Closed-form numpy restatement of the per-shift counts -- TEST INFRASTRUCTURE ONLY.

An algorithmically independent second checker (boolean arrays instead of packed
words) for small cases.  Same import rule as cc_oracle.py: tests/, smoke() and
bench.py's CPU legs only.  Parity status: pinned through tests/test_oracle_golden.py.

Formulas (verified there against the compiled reference):
with zero-padded 1-based boolean arrays F, R, M of the reference's length
``chrom_len + read_len + max_shift + 100`` (mscc.pyx:134),

    ncc[k]  = sum_j F[j] R[j+k]                      mscc.pyx:314
    D_k[j]  = M[j] M[j + read_len-1-k]               mscc.pyx:283-284, :291, :307-310
    dmp[k]  = sum_j D_k[j]                           mscc.pyx:293-298
    mfwd[k] = sum_j F[j] D_k[j]                      mscc.pyx:300, :303
    mrev[k] = sum_j R[j+k] D_k[j]                    mscc.pyx:301, :304
    mscc[k] = sum_j F[j] R[j+k] D_k[j]               mscc.pyx:305
"""
import numpy as np

EXTRA_ALLOCATE_SIZE = 100


def _shifted(a, off):
    """b[j] = a[j+off], zero where j+off falls outside the array."""
    n = a.size
    b = np.zeros_like(a)
    if off >= 0:
        if off < n:
            b[:n - off] = a[off:]
    else:
        if -off < n:
            b[-off:] = a[:n + off]
    return b


def chrom_counts(chrom_len, read_len, max_shift, fwd_pos, rev_pos,
                 map_begins=None, map_ends=None, skip_ncc=False, lo=0, hi=None):
    nbits = chrom_len + read_len + max_shift + EXTRA_ALLOCATE_SIZE
    if hi is None:
        hi = nbits
    F = np.zeros(nbits, dtype=bool)
    R = np.zeros(nbits, dtype=bool)
    F[np.asarray(fwd_pos, dtype=np.int64)] = True
    R[np.asarray(rev_pos, dtype=np.int64)] = True
    M = None
    if map_begins is not None:
        M = np.zeros(nbits, dtype=bool)
        for b, e in zip(map_begins, map_ends):
            M[int(b) + 1:int(e) + 1] = True
    sel = np.zeros(nbits, dtype=bool)
    sel[lo:hi] = True
    S = max_shift + 1
    out = {"fwd_total": int((F & sel).sum()), "rev_total": int((R & sel).sum())}
    ncc = np.zeros(S, dtype=np.int64)
    dmp, mf, mr, ms = (np.zeros(S, dtype=np.int64) for _ in range(4))
    for k in range(S):
        Rk = _shifted(R, k)
        ncc[k] = (F & Rk & sel).sum()
        if M is not None:
            D = M & _shifted(M, read_len - 1 - k) & sel
            dmp[k] = D.sum()
            mf[k] = (F & D).sum()
            mr[k] = (Rk & D).sum()
            ms[k] = (F & Rk & D).sum()
    if not skip_ncc:
        out["ncc"] = ncc
    if M is not None:
        out.update(dmp=dmp, mfwd=mf, mrev=mr, mscc=ms)
    return out
