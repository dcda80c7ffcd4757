"""Pure-Python model of the LIST path's arithmetic (pymasc_b200/csrc/pmb_lists.cuh) -- TEST
INFRASTRUCTURE ONLY, like everything under oracle/.

The list path never builds bit arrays: every count of mscc.pyx:288-317 is a sum over PAIRS of
sorted list entries (reads x reads, reads x mappable runs, runs x runs) that fall within
max_shift of each other.  This file restates, thread for thread, what one CUDA thread of each
list kernel adds to the histograms -- with the kernels' own devices where they carry risk:

  * 32-bit unsigned arithmetic that wraps (`u32`), as the kernels compute: the run a shift
    can meet as (k1, span) with `k - k1 <= span`, "no run" as k1 = NONE, span = 0; the run-edge
    kernel's distances modulo 2^32 tested against max_shift as unsigned numbers;
  * the searches as the kernels make them: a warp-uniform cursor, six probes at fixed offsets
    into the 63 entries behind it, the gallop + binary search a vote falls back to;
  * the forward kernel's packed pair histogram: ncc and mscc of a shift in the low / high half
    of ONE 32-bit bin (kept modulo 2^32 here, so an overflow of a half would show), flushed
    every `pair_cap` reads; of equal reverse reads the LAST is counted;
  * groups of 32 entries with the cursor handed on by the last lane, the checked and the
    unchecked pair loop chosen by the group's vote.

so that the formulas can be checked against cc_oracle.c on the CPU (tests/test_list_model.py)
without a GPU.  What it does not model is scheduling: one warp walks a whole list here.

Inputs follow include/pymasc_b200.h: sorted forward 5' positions, sorted reverse 5' ends,
sorted disjoint BigWig intervals (bits begin+1 .. end), counted range [lo, hi).
"""
import bisect

import numpy as np

M32 = 0xFFFFFFFF
NONE = 0xFFFFFFFF
PAIR_CAP = 65535            # LIST_PAIR_CAP


def u32(x):
    return x & M32


# ---------------------------------------------------------------------------- searches
def gallop(a, n, cur, key_max):
    step = 48
    while True:
        if n - cur <= step:
            return n
        if a[cur + step] >= key_max:
            return cur + step
        cur += step
        step <<= 1


def lower_bound_in(a, lo, hi, key):
    while lo < hi:
        mid = (lo + hi) >> 1
        if a[mid] < key:
            lo = mid + 1
        else:
            hi = mid
    return lo


def cursor_window(a, n, cur, key):
    """Offset of the first entry >= key among the 63 after the cursor; 63: all of them are
    below (or the list ends inside the window)."""
    off = 63
    if cur + 63 <= n:
        off = 0
        s = 32
        while s:
            if a[cur + off + (s - 1)] < key:
                off += s
            s >>= 1
    return off


def group_lower_bounds(a, n, cur, keys, key_max):
    """What the lanes of one group get from cursor_window + cursor_resolve (the vote is the group's)."""
    offs = [cursor_window(a, n, cur, k) for k in keys]
    if any(o == 63 for o in offs):
        hi = gallop(a, n, cur, key_max)
        return [lower_bound_in(a, cur, hi, k) for k in keys]
    return [cur + o for o in offs]


def _groups(n):
    for base in range(0, n, 32):
        yield base, min(32, n - base)


# ---------------------------------------------------------------------------- forward reads
def forward_reads(fwd, rev, begins, ends, rl, D, lo, hi, nbits, ncc=True, mscc=True, pair_cap=PAIR_CAP):
    """list_fwd_kernel: one thread per forward read.  Returns ncc, mscc (pair histograms),
    fplus/fminus (difference histogram of mfwd), fwd_total."""
    S2, DUMP = D + 2, D + 1                      # bins 0..D and the dump bin D+1
    acc_ncc, acc_mscc = [0] * S2, [0] * S2      # the 64-bit global accumulators
    pair, fp, fm = [0] * S2, [0] * S2, [0] * S2
    fwd = [int(x) for x in fwd]
    rev = [int(x) for x in rev]
    begins = [int(x) for x in begins] if begins is not None else []
    ends = [int(x) for x in ends] if ends is not None else []
    nR, nI = len(rev), len(begins)
    job_mscc = bool(mscc and nI)
    rl1 = rl - 1
    r_last = rev[nR - 1] if nR else 0
    total = 0

    def flush():
        for k in range(D + 1):
            v = pair[k]
            if mscc:
                acc_ncc[k] += v & 0xFFFF
                acc_mscc[k] += v >> 16
            else:
                acc_ncc[k] += v
            pair[k] = 0

    def segment(w0, w1):
        nonlocal total
        JR = bisect.bisect_left(rev, fwd[w0])                        # warp_lower_bound
        CI = bisect.bisect_left(begins, fwd[w0] + rl1) if job_mscc else 0
        for base in range(w0, w1, 32):
            n_act = min(32, w1 - base)
            ps = fwd[base:base + n_act]
            p_max = ps[-1]
            valid = [(base + l == 0 or p != fwd[base + l - 1]) and lo <= p < hi for l, p in enumerate(ps)]
            total += sum(valid)
            js = group_lower_bounds(rev, nR, JR, ps, p_max)
            JR = js[-1]
            cs = [None] * n_act
            if job_mscc:
                c1 = group_lower_bounds(begins, nI, CI, [u32(p + rl1) for p in ps], u32(p_max + rl1))
                CI = c1[-1]
                cs = [x - 1 for x in c1]
            fast = all((not v) or (j < nR and r_last > p and r_last - p > D) for v, j, p in zip(valid, js, ps))
            for p, ok, j, c in zip(ps, valid, js, cs):
                if not ok:
                    continue
                x_top = u32(p + rl1)
                mp = False
                if job_mscc:
                    cp = c
                    while cp >= 0 and begins[cp] >= p:
                        cp -= 1
                    mp = cp >= 0 and p <= ends[cp]                # M[p]
                k1, span, k2w = NONE, 0, NONE

                def load_run(c):
                    b, e = begins[c], ends[c]
                    k2 = u32(x_top - b - 1)
                    r1 = u32(x_top - min(e, x_top))
                    if e > b:
                        fp[min(r1, DUMP)] += 1
                        fm[min(u32(k2 + 1), DUMP)] += 1
                        return r1, u32(k2 - r1), (k2 if c > 0 else NONE)
                    return NONE, 0, (k2 if c > 0 else NONE)

                if mscc and mp and c >= 0:
                    k1, span, k2w = load_run(c)
                else:
                    mp = False
                # the pair loop; `checked`: the list may end inside this group's windows
                checked = not fast
                more = (not checked) or j < nR
                q = rev[j] if more else 0
                k = u32(q - p)
                while more and k <= D:
                    j += 1
                    last = checked and j >= nR
                    qn = u32(~q) if last else rev[j]              # (unchecked: the vote says rev[j] exists)
                    once = q != qn
                    if mscc:
                        while k > k2w:
                            c -= 1
                            k1, span, k2w = load_run(c)
                        pair[k] = u32(pair[k] + ((0x10001 if u32(k - k1) <= span else 1) if once else 0))
                    else:
                        pair[k if once else DUMP] += 1
                    q = qn
                    k = u32(q - p)
                    more = not last
                if mscc and mp:
                    while k2w < D:
                        c -= 1
                        k1, span, k2w = load_run(c)

    n = len(fwd)
    cap = max(1, min(pair_cap, PAIR_CAP)) if mscc else n + 1
    pos = 0
    while pos < n:                                  # list_segments: a segment is cut after `cap` reads
        seg_hi = min(n, pos + cap)
        segment(pos, seg_hi)
        flush()
        pos = seg_hi
    if not ncc:
        acc_ncc = [0] * S2
    return acc_ncc, acc_mscc, fp, fm, total


# ---------------------------------------------------------------------------- reverse reads
def reverse_reads(rev, begins, ends, rl, D, lo, hi, nbits, mscc=True):
    """list_rev_kernel: one thread per reverse read.  Returns rplus/rminus (difference
    histogram of mrev) and rev_total."""
    S2 = D + 2
    rp, rm = [0] * S2, [0] * S2
    begins = [int(x) for x in begins] if begins is not None else []
    ends = [int(x) for x in ends] if ends is not None else []
    rev = [int(x) for x in rev]
    nI = len(begins)
    job_mscc = bool(mscc and nI)
    rl1 = rl - 1
    total = 0
    if not rev:
        return rp, rm, 0
    CI = bisect.bisect_left(begins, rev[0] + rl1) if job_mscc else 0
    for base, n_act in _groups(len(rev)):
        qs = rev[base:base + n_act]
        valid = [(base + l == 0 or q != rev[base + l - 1]) and q < nbits for l, q in enumerate(qs)]
        total += sum(1 for v, q in zip(valid, qs) if v and lo <= q < hi)
        if not job_mscc:
            continue
        c1 = group_lower_bounds(begins, nI, CI, [u32(q + rl1) for q in qs], u32(qs[-1] + rl1))
        CI = c1[-1]
        for q, ok, cb1 in zip(qs, valid, c1):
            t = u32(q + rl1)
            ka = min(u32(q - hi + 1), D + 1) if q >= hi else 0
            kb = min(u32(q - lo), D)
            cB = cb1 - 1
            cA = cB
            while cA >= 0 and begins[cA] >= q:
                cA -= 1
            if not ok or q < lo or ka > kb or cA < 0:
                continue

            def load_a(c):
                b, e = begins[c], ends[c]
                return u32(q - min(e, q)), u32(q - b - 1)

            def load_b(c):
                b, e = begins[c], ends[c]
                return u32(t - min(e, t) + 1) >> 1, u32(t - b - 1) >> 1

            a1, a2 = load_a(cA)
            b1, b2 = load_b(cB)
            while True:
                if a1 > kb or b1 > kb:
                    break
                s, e = max(a1, b1, ka), min(a2, b2, kb)
                if s <= e:                      # (else: both updates go to the dump bin)
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


# ---------------------------------------------------------------------------- run edges
def edge_in_window(edge, base, D):
    """The run-edge kernel's test: distance modulo 2^32 against max_shift as unsigned numbers.
    Returns the bin or None.  Exact while chromosomes keep 2 (max_shift + read_len) clear of
    2^32 bits (pmb_api.cu, list_ok): tests/test_list_model.py checks it against plain integers."""
    d = u32(edge - base)
    return d if d <= D else None


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
        w_hi = u32(t + off + D)
        key = max(s + off - 1, 0)
        J = i                                   # walked to, from the thread's own run
        while J > 0 and ends[J - 1] >= key:
            J -= 1
        if J == i:
            while J < n and ends[J] < key:
                J += 1
        base0, base1 = u32(s + off), u32(t + off)
        while J < n:
            bj, ej = begins[J], ends[J]
            if u32(bj + 1) > w_hi:
                break
            J += 1
            if ej <= bj:
                continue
            for edge, sign in ((bj + 1, +1), (ej + 1, -1)):
                k = edge_in_window(u32(edge), base0, D)     # start . (start | end)
                if k is not None:
                    (plus if sign > 0 else minus)[k] += 1
                k = edge_in_window(u32(edge), base1, D)     # end . (start | end)
                if k is not None:
                    (minus if sign > 0 else plus)[k] += 1
        # direct lags: overlap of this (clipped) run with the runs at and above it, moved down
        A_lo, A_hi = s, t - 1
        kk = i
        while kk < n:
            B_lo, B_hi = u32(begins[kk] + 1), ends[kk]
            if B_lo > u32(A_hi + Lmax):
                break
            kk += 1
            if B_hi < B_lo:
                continue
            for L in (L0, L1):
                if B_hi >= L:
                    top, bot = min(A_hi, B_hi - L), max(A_lo, B_lo - L if B_lo > L else 0)
                    if top >= bot:
                        if L == L0:
                            d0 += top - bot + 1
                        else:
                            d1 += top - bot + 1
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


def chrom_counts(chrom_len, rl, D, fwd, rev, begins=None, ends=None, skip_ncc=False, lo=0, hi=None,
                 pair_cap=PAIR_CAP):
    """Same contract as oracle.cc_oracle.chrom_counts, computed the list path's way."""
    nbits = chrom_len + rl + D + 100
    hi = nbits if hi is None else min(hi, nbits)
    lo = min(lo, hi)
    mscc = begins is not None
    h_ncc, h_mscc, fp, fm, ft = forward_reads(fwd, rev, begins, ends, rl, D, lo, hi, nbits,
                                              ncc=not skip_ncc, mscc=mscc, pair_cap=pair_cap)
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
