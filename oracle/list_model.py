"""Pure-Python model of the LIST path's arithmetic (pymasc_b200/csrc/pmb_lists.cuh) -- TEST
INFRASTRUCTURE ONLY, like everything under oracle/.

The list path never builds bit arrays: every count of mscc.pyx:288-317 is a sum over PAIRS of
sorted list entries (reads x reads, reads x mappable runs, runs x runs) that fall within
max_shift of each other.  This file restates, thread for thread, what one CUDA thread of each
list kernel adds to the histograms, so that the formulas (run -> shift ranges, clamps, the
difference histograms and their prefix sums, the dump bin) can be checked against cc_oracle.c
on the CPU (tests/test_list_model.py) before the kernels ever see a GPU.

What the model does not restate is how the kernels are shaped for the machine: of equal reverse
reads the kernel counts the LAST (the model the first -- a bit is set once either way), ncc and
mscc of a shift share one 32-bit bin there, and its searches look at a window after a cursor.

Inputs follow include/pymasc_b200.h: sorted forward 5' positions, sorted reverse 5' ends,
sorted disjoint BigWig intervals (bits begin+1 .. end), counted range [lo, hi).
"""
import bisect

import numpy as np


def _dedup_sorted(a):
    a = [int(x) for x in a]
    return [x for i, x in enumerate(a) if i == 0 or x != a[i - 1]]


def forward_reads(fwd, rev, begins, ends, rl, D, lo, hi, nbits, ncc=True, mscc=True):
    """list_fwd_kernel: one thread per forward read.  Returns ncc, mscc (pair histograms),
    fplus/fminus (difference histogram of mfwd), fwd_total."""
    S2 = D + 2                                   # bins 0..D and the dump bin D+1
    h_ncc, h_mscc, fp, fm = [0] * S2, [0] * S2, [0] * S2, [0] * S2
    rev = [int(x) for x in rev]
    begins = [int(x) for x in begins] if begins is not None else []
    ends = [int(x) for x in ends] if ends is not None else []
    total = 0
    fwd = [int(x) for x in fwd]
    for i, p in enumerate(fwd):
        if i and p == fwd[i - 1]:
            continue
        if not (lo <= p < hi) or p >= nbits:
            continue
        total += 1
        jr = bisect.bisect_left(rev, p)
        x_top = p + rl - 1
        c = bisect.bisect_left(begins, x_top) - 1      # last interval with begin < x_top
        cp = c
        while cp >= 0 and begins[cp] >= p:
            cp -= 1
        mp = mscc and cp >= 0 and p <= ends[cp]
        k1 = k2 = 1 << 40                               # "no run"

        def load_run(c):
            b, e = begins[c], ends[c]
            r2 = x_top - b - 1
            r1 = x_top - min(e, x_top)
            if e > b:
                fp[min(r1, D + 1)] += 1
                fm[min(r2 + 1, D + 1)] += 1
            return r1, r2

        if mp and c >= 0:
            k1, k2 = load_run(c)
        qprev = None
        j = jr
        while j < len(rev):
            q = rev[j]
            j += 1
            k = q - p
            if k > D:
                break
            if q == qprev:
                continue
            qprev = q
            if ncc:
                h_ncc[k] += 1
            if mp:
                while k > k2 and c > 0:
                    c -= 1
                    k1, k2 = load_run(c)
                if k1 <= k <= k2:
                    h_mscc[k] += 1
        if mp:
            while k2 < D and c > 0:
                c -= 1
                k1, k2 = load_run(c)
    return h_ncc, h_mscc, fp, fm, total


def reverse_reads(rev, begins, ends, rl, D, lo, hi, nbits, mscc=True):
    """list_rev_kernel: one thread per reverse read.  Returns rplus/rminus (difference
    histogram of mrev) and rev_total."""
    S2 = D + 2
    rp, rm = [0] * S2, [0] * S2
    begins = [int(x) for x in begins] if begins is not None else []
    ends = [int(x) for x in ends] if ends is not None else []
    rev = [int(x) for x in rev]
    total = 0
    for j, q in enumerate(rev):
        if j and q == rev[j - 1]:
            continue
        if q >= nbits:
            continue
        if lo <= q < hi:
            total += 1
        if not mscc:
            continue
        ka, kb = max(0, q - hi + 1), min(D, q - lo)
        if ka > kb:
            continue
        t = q + rl - 1
        cB = bisect.bisect_left(begins, t) - 1
        cA = cB
        while cA >= 0 and begins[cA] >= q:
            cA -= 1
        if cA < 0 or cB < 0:
            continue

        def load_a(c):
            b, e = begins[c], ends[c]
            return q - min(e, q), q - b - 1

        def load_b(c):
            b, e = begins[c], ends[c]
            return (t - min(e, t) + 1) >> 1, (t - b - 1) >> 1

        a1, a2 = load_a(cA)
        b1, b2 = load_b(cB)
        while True:
            if a1 > kb or b1 > kb:
                break
            s, e = max(a1, b1, ka), min(a2, b2, kb)
            if s <= e:
                rp[s] += 1
                rm[e + 1] += 1
            if min(a2, b2) >= kb:
                break
            if a2 <= b2:
                if cA == 0:
                    break
                cA -= 1
                a1, a2 = load_a(cA)
            else:
                if cB == 0:
                    break
                cB -= 1
                b1, b2 = load_b(cB)
    return rp, rm, total


def run_edges(begins, ends, rl, D, lo, hi, autocorr=False):
    """list_edge_kernel: one thread per mappable interval.  Returns the edge products
    plus/minus (window-bit indexed, as finalize wants them) and the two direct lags."""
    S2 = D + 2
    plus, minus = [0] * S2, [0] * S2
    begins = [int(x) for x in begins]
    ends = [int(x) for x in ends]
    n = len(begins)
    off = 0 if autocorr else rl - 1 - D
    L0, L1 = (0, 1) if autocorr else (rl - 1, rl - 2)
    Lmax = max(L0, L1)
    d0 = d1 = 0
    for i in range(n):
        b, e = begins[i], ends[i]
        s, t = max(b + 1, lo), min(e + 1, hi)
        if s >= t:
            continue
        # run edges of M inside the windows of this run's two edges
        w_lo, w_hi = s + off, t + off + D
        jl = bisect.bisect_left(ends, w_lo - 1)
        J = jl
        while J < n and begins[J] + 1 <= w_hi:
            bj, ej = begins[J], ends[J]
            J += 1
            if ej <= bj:
                continue
            for v, v_is_end in ((bj + 1, False), (ej + 1, True)):
                d = v - (s + off)
                if 0 <= d <= D:
                    (minus if v_is_end else plus)[d] += 1          # start . (start | end)
                d = v - (t + off)
                if 0 <= d <= D:
                    (plus if v_is_end else minus)[d] += 1          # end . (end | start)
        # direct lags: overlap of this (clipped) run with the runs at and above it, moved down
        A_lo, A_hi = s, t - 1
        kk = i
        while kk < n:
            B_lo, B_hi = begins[kk] + 1, ends[kk]
            kk += 1
            if B_lo - Lmax > A_hi:
                break
            if B_hi < B_lo:
                continue
            o0 = min(A_hi, B_hi - L0) - max(A_lo, B_lo - L0) + 1
            if o0 > 0:
                d0 += o0
            o1 = min(A_hi, B_hi - L1) - max(A_lo, B_lo - L1) + 1
            if o1 > 0:
                d1 += o1
    return plus, minus, d0, d1


def finalize_dmp(plus, minus, a0, a1, D, autocorr=False):
    """finalize: dmp from the two direct lags and C[b] = plus[b] - minus[b] (double prefix sum)."""
    if D == 0:
        return [a0]
    out = [0] * (D + 1)
    out[0], out[1] = a0, a1 - a0
    for k in range(2, D + 1):
        b = (k - 1) if autocorr else D - (k - 1)
        out[k] = -(plus[b] - minus[b])
    for k in range(2, D + 1):           # first differences d[1..D]
        out[k] += out[k - 1]
    for k in range(1, D + 1):
        out[k] += out[k - 1]
    return out


def prefix(plus, minus, D):
    out, run = [], 0
    for k in range(D + 1):
        run += plus[k] - minus[k]
        out.append(run)
    return out


def chrom_counts(chrom_len, rl, D, fwd, rev, begins=None, ends=None, skip_ncc=False, lo=0, hi=None):
    """Same contract as oracle.cc_oracle.chrom_counts, computed the list path's way."""
    nbits = chrom_len + rl + D + 100
    hi = nbits if hi is None else min(hi, nbits)
    lo = min(lo, hi)
    mscc = begins is not None
    h_ncc, h_mscc, fp, fm, ft = forward_reads(fwd, rev, begins, ends, rl, D, lo, hi, nbits,
                                              ncc=not skip_ncc, mscc=mscc)
    rp, rm, rt = reverse_reads(rev, begins, ends, rl, D, lo, hi, nbits, mscc=mscc)
    out = {"fwd_total": ft, "rev_total": rt}
    if not skip_ncc:
        out["ncc"] = np.array(h_ncc[:D + 1], np.int64)
    if mscc:
        plus, minus, a0, a1 = run_edges(begins, ends, rl, D, lo, hi)
        out["dmp"] = np.array(finalize_dmp(plus, minus, a0, a1, D), np.int64)
        out["mfwd"] = np.array(prefix(fp, fm, D), np.int64)
        out["mrev"] = np.array(prefix(rp, rm, D), np.int64)
        out["mscc"] = np.array(h_mscc[:D + 1], np.int64)
    return out


def mappable_autocorr(chrom_len, rl, D, begins, ends, lo=0, hi=None):
    nbits = chrom_len + rl + D + 100
    hi = nbits if hi is None else min(hi, nbits)
    plus, minus, a0, a1 = run_edges(begins, ends, rl, D, lo, hi, autocorr=True)
    return np.array(finalize_dmp(plus, minus, a0, a1, D, autocorr=True), np.int64)
