// pmb_lists.cuh -- the LIST path (sm_100a): the same per-shift counts as the shift-window and
// event-window kernels, computed straight from the SORTED position and interval lists the host
// hands over.  No bit array is ever built.
//
// Every count of the reference's loop (PyMaSC/core/bitarray/mscc.pyx:288-317) is a sum over
// pairs of list entries that lie within max_shift of each other:
//
//   ncc[k]  = #{(p, q): q - p = k}                       p forward 5' end, q reverse 5' end   (:314)
//   mscc[k] = the same pairs with M[p] and M[p+rl-1-k]                                        (:305)
//   mfwd[k] = sum_p M[p] M[p+rl-1-k]    a step function of k per read: +1 where a mappable
//             run enters the window, -1 where it leaves -> a DIFFERENCE histogram, prefix-
//             summed once at the end                                                          (:303)
//   mrev[k] = sum_q M[q-k] M[q+rl-1-2k] the product of two step functions (runs met at stride
//             1 and at stride 2): merge the two run lists, again +1 / -1 per piece            (:304)
//   dmp[k]  = A[rl-1-k], A[s] = sum_j Mc[j] M[j+s]: its second difference is a sum over pairs
//             of run edges; two directly counted lags anchor it (as on the event path)        (:291-298)
//
// With 20 M reads on hg38 a forward read sees ~3 reverse reads and ~1 run edge inside its 1000
// shifts, so a read costs a few histogram updates instead of a 1000-bit window per counter.
// Histograms live in shared memory (ATOMS: ~11 updates/clk/SM measured on B200,
// profiles/r02_atom_probe.jsonl -- the cost is per instruction, whatever the bins), one copy
// per CTA, flushed to the 64-bit global counters when the CTA changes chromosome.  The inner
// loops are shaped against the SASS: shared-space addresses converted once,
// conditions that select the value added or a dump bin instead of branching around an update,
// ncc and mscc of a shift in one 32-bit bin.
//
// Mapping.  One THREAD per list entry.  A CTA owns a contiguous range of the batch's entries,
// cut into per-chromosome segments; inside a segment every warp owns a contiguous sub-range and
// walks it 32 entries at a time with warp-uniform cursors into the other lists: each entry's own
// start is searched in the 63 entries after the cursor (six probes at fixed offsets; a gallop and
// a binary search only where a vote says the window may be too short).  Warps meet only where the
// CTA changes chromosome.
//
// Preconditions, verified on the device (flag -> the host re-runs the batch on the event path):
// reads sorted; intervals sorted and disjoint.  Exact integer arithmetic throughout; the
// thread-for-thread CPU restatement of this file is oracle/list_model.py.
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>
#include <type_traits>

#include "pmb_events.cuh"   // EVF_*, C_*, block_scan_inplace

namespace pmb {

// Accumulators.  The list kernels ADD into per-job accumulators, the finalize kernel turns them
// into the job's output vectors with plain stores.  There are two copies of every accumulator
// (and of the flags), used by alternate executes of a batch: the finalize of execute k, once it
// has read copy k & 1, clears copy (k+1) & 1 -- dirty since execute k-1, which nobody reads any
// more -- so a step needs no zeroing pass and no kernel ever clears what another still reads.
//   acc  [LA_N][stride] | [2] strand totals        per job
//   aux  [2][stride] run-edge products | [2] lags  per job that OWNS its track (followers point at it)
enum { LA_NCC = 0, LA_MSCC = 1, LA_FP = 2, LA_FM = 3, LA_RP = 4, LA_RM = 5, LA_N = 6 };

struct ListJob {
    const uint32_t* fwd;
    const uint32_t* rev;
    const uint32_t* begins;         // the owner's intervals for a job that shares its track
    const uint32_t* ends;
    uint32_t n_fwd, n_rev, n_iv, flags;
    unsigned long long* counters;   // out: [5][stride]
    unsigned long long* totals;     // out: [2]
    unsigned long long* acc;        // copy 0 of this job's accumulators
    unsigned long long* aux;        // copy 0 of the (owner's) run-edge products and lags
    long long lo, hi;               // counted range, chromosome bits
    unsigned long long nbits;
};

struct ListParams {
    const ListJob* jobs;
    const unsigned long long* prefix;   // [n_jobs + 1]: entries of this kernel's list before each job
    uint32_t job_lo, job_hi;            // this launch covers these jobs (everything, or one H2D group)
    uint32_t max_shift, read_len, stride;
    uint32_t pair_cap;                  // forward kernel: reads between two flushes of the packed pair histogram (<= LIST_PAIR_CAP)
    int* flags;                         // this execute's: [0] position beyond the array, [1] input not sorted / disjoint
    unsigned long long acc_off;         // words from copy 0 to this execute's copy of the accumulators
    // finalize only
    unsigned long long acc_other;       // ... to the copy the NEXT execute uses (cleared here); == acc_off: clear nothing
    int* flags_other;                   // the next execute's flags (cleared here), or null
    unsigned long long* status_tail;    // [2] status words behind the counter block (pmb_allreduce), or null
};

#ifndef PMB_LIST_THREADS
#define PMB_LIST_THREADS 512
#endif
#ifndef PMB_LIST_MIN_BLOCKS
#define PMB_LIST_MIN_BLOCKS 2
#endif
constexpr int LIST_THREADS = PMB_LIST_THREADS;
constexpr int LIST_MIN_BLOCKS = PMB_LIST_MIN_BLOCKS;      // resident CTAs per SM the register budget is cut for
constexpr uint32_t LIST_NONE = 0xFFFFFFFFu;

// (PMB_CHECK: pmb_window.cuh -- bounds checks of every list index and histogram bin in a -DPMB_CHECKED build)

__host__ __device__ constexpr size_t list_smem_bytes(uint32_t max_shift, int n_hist)
{
    return (size_t)n_hist * (max_shift + 2) * 4;      // bins 0 .. max_shift and a dump bin
}
__host__ __device__ constexpr int list_fwd_hists(bool mscc) { return mscc ? 3 : 1; }   // pairs | d(mfwd) plus, minus
constexpr unsigned LIST_PAIR_CAP = 65535;             // forward reads between two flushes of the packed pair histogram

// ---------------------------------------------------------------- searches
// number of a[0..n) that are < key: warp-cooperative 33-ary search (one probe per lane and round)
__device__ __forceinline__ uint32_t warp_lower_bound(const uint32_t* __restrict__ a, uint32_t n, uint32_t key, int lane)
{
    uint32_t lo = 0, hi = n;                    // the answer lies in [lo, hi]
    while (hi - lo >= 33) {
        const uint32_t width = hi - lo;
        const uint32_t pi = lo + (uint32_t)(((uint64_t)width * (lane + 1)) / 33);
        const uint32_t cnt = __popc(__ballot_sync(0xFFFFFFFFu, __ldg(a + pi) < key));     // a prefix of the lanes
        const uint32_t nlo = cnt ? lo + (uint32_t)(((uint64_t)width * cnt) / 33) + 1 : lo;
        const uint32_t nhi = cnt < 32 ? lo + (uint32_t)(((uint64_t)width * (cnt + 1)) / 33) : hi;
        lo = nlo; hi = nhi;
    }
    const uint32_t pi = lo + lane;
    return lo + __popc(__ballot_sync(0xFFFFFFFFu, pi < hi && __ldg(a + pi) < key));
}

// first index in [lo, hi] with a[index] >= key, given a[< lo] < key and (hi == n or a[hi] >= key)
__device__ __forceinline__ uint32_t lower_bound_in(const uint32_t* __restrict__ a, uint32_t lo, uint32_t hi, uint32_t key)
{
    while (lo < hi) {
        const uint32_t mid = (lo + hi) >> 1;
        if (__ldg(a + mid) < key) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// warp-uniform: from a cursor with a[< cur] < every key of the group, an index `hi` with
// hi == n or a[hi] >= key_max (doubling steps; one or two probes for evenly spread lists)
__device__ __forceinline__ uint32_t gallop(const uint32_t* __restrict__ a, uint32_t n, uint32_t cur, uint32_t key_max)
{
    uint32_t step = 48;
    for (;;) {
        if (n - cur <= step) return n;
        if (__ldg(a + cur + step) >= key_max) return cur + step;
        cur += step;                            // everything up to here is below every later key, too
        step <<= 1;
    }
}

// The search every group makes: first index >= cur with a[index] >= key, given a[< cur] < key and
// a warp-uniform cursor.  The answer nearly always lies within the next 63 entries (a group of 32
// reads meets about as many entries of the other list), so those are searched without a loop:
// six probes at fixed offsets from one base address.  If any lane finds all 63 below its key --
// or the list ends inside the window -- the warp bounds the range by galloping and searches that.
// (key_max: the largest key of the group.  Unsorted input gives wrong but in-bounds answers.)
// (in two steps, so that a caller with two independent searches can issue both probe chains
// before either vote: the probes are dependent L1 hits, the chains overlap)
__device__ __forceinline__ uint32_t cursor_window(const uint32_t* __restrict__ a, uint32_t n, uint32_t cur, uint32_t key)
{
    uint32_t off = 63;
    if (cur + 63 <= n) {
        const uint32_t* __restrict__ b = a + cur;
        off = 0;
#pragma unroll
        for (uint32_t s = 32; s; s >>= 1)
            if (__ldg(b + off + (s - 1)) < key) off += s;
    }
    return off;
}
__device__ __forceinline__ uint32_t cursor_resolve(const uint32_t* __restrict__ a, uint32_t n, uint32_t cur,
                                                   uint32_t key, uint32_t key_max, uint32_t off)
{
    if (__any_sync(0xFFFFFFFFu, off == 63)) {
        const uint32_t hi = gallop(a, n, cur, key_max);
        return lower_bound_in(a, cur, hi, key);
    }
    return cur + off;
}
__device__ __forceinline__ uint32_t cursor_lower_bound(const uint32_t* __restrict__ a, uint32_t n, uint32_t cur,
                                                       uint32_t key, uint32_t key_max)
{
    return cursor_resolve(a, n, cur, key, key_max, cursor_window(a, n, cur, key));
}

// ---------------------------------------------------------------- the CTA / segment / warp skeleton
// body(job, J, w0, w1): called by every warp with its sub-range [w0, w1) of the job's entries
// flush(job, J): called by the whole CTA (between barriers) when a segment ends
// seg_cap: a segment never holds more entries than this (the forward kernel's packed pair
// histogram counts in 16 bits between flushes)
template <class Body, class Flush>
__device__ __forceinline__ void list_segments(const ListParams& P, int warp, int n_warps, Body&& body, Flush&& flush,
                                              unsigned long long seg_cap = ~0ull)
{
    const unsigned long long g0 = P.prefix[P.job_lo], g1 = P.prefix[P.job_hi];
    const unsigned long long N = g1 - g0;
    const unsigned long long c_lo = g0 + (N * blockIdx.x) / gridDim.x;
    const unsigned long long c_hi = g0 + (N * (blockIdx.x + 1)) / gridDim.x;
    if (c_lo >= c_hi) return;
    uint32_t job = P.job_lo;
    while (P.prefix[job + 1] <= c_lo) job++;
    for (unsigned long long pos = c_lo; pos < c_hi;) {
        const unsigned long long jb = P.prefix[job], je = P.prefix[job + 1];
        if (je <= pos) { ++job; continue; }                     // a job without entries, or done with
        unsigned long long seg_hi = c_hi < je ? c_hi : je;
        if (seg_hi - pos > seg_cap) seg_hi = pos + seg_cap;
        const ListJob& J = P.jobs[job];
        const uint32_t s0 = (uint32_t)(pos - jb), s1 = (uint32_t)(seg_hi - jb);
        PMB_CHECK(job < P.job_hi && s0 < s1 && jb + s1 <= je, P.flags);
        const uint32_t len = s1 - s0;
        const uint32_t chunk = (((len + n_warps - 1) / n_warps) + 31u) & ~31u;
        const unsigned long long w0 = (unsigned long long)s0 + (unsigned long long)warp * chunk;
        if (w0 < s1) body(job, J, (uint32_t)w0, (uint32_t)min((unsigned long long)s1, w0 + chunk));
        __syncthreads();
        flush(job, J);
        __syncthreads();
        pos = seg_hi;
    }
}

// add the non-zero bins of a shared-memory histogram to 64-bit global counters and clear them
__device__ __forceinline__ void flush_hist(uint32_t* h, uint32_t n_bins, unsigned long long* dst)
{
    for (uint32_t i = threadIdx.x; i < n_bins; i += blockDim.x) {
        const uint32_t v = h[i];
        if (v) { atomicAdd(dst + i, (unsigned long long)v); h[i] = 0; }
    }
}

// Shared-memory histogram updates by 32-bit shared-space address.  With generic pointers the
// compiler rebuilt the address of every update inside the pair loop (S2UR SR_CgaCtaId, ULEA, LEA:
// the shared window of the CTA) and branched around each atomicAdd; the loop was 36 SASS
// instructions per pair.
__device__ __forceinline__ uint32_t smem_addr(const void* p)
{
    uint32_t a;     // (volatile: the compiler may not redo the conversion where the address is used -- it did, per pair)
    asm volatile("{ .reg .u64 t;\n\tcvta.to.shared.u64 t, %1;\n\tcvt.u32.u64 %0, t; }" : "=r"(a) : "l"(p));
    return a;
}
// ++[addr]; callers that count conditionally pass the address of a dump bin instead of branching
// around the update (ptxas does not predicate ATOMS: every condition became BSSY / BRA / BSYNC,
// and an update costs one issue slot whatever the number of lanes that take part)
__device__ __forceinline__ void smem_inc(uint32_t addr)
{
    asm volatile("red.shared.add.u32 [%0], 1;" :: "r"(addr) : "memory");
}
__device__ __forceinline__ void smem_add(uint32_t addr, uint32_t v)
{
    asm volatile("red.shared.add.u32 [%0], %1;" :: "r"(addr), "r"(v) : "memory");
}
// ---------------------------------------------------------------- forward reads: ncc, mscc, d(mfwd)
template <bool NCC, bool MSCC>
__global__ void __launch_bounds__(LIST_THREADS, LIST_MIN_BLOCKS) list_fwd_kernel(const ListParams P)
{
    extern __shared__ uint32_t lsm[];
    const uint32_t D = P.max_shift, S2 = D + 2, DUMP = D + 1;
    // MSCC: ncc and mscc of a shift share one 32-bit bin (low / high half), so a pair costs ONE
    // update; a read adds at most 1 to a bin, and a segment is cut after LIST_PAIR_CAP reads
    uint32_t* h_pair = lsm;
    uint32_t* h_fp = lsm + S2;
    uint32_t* h_fm = lsm + 2 * S2;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, n_warps = blockDim.x >> 5;
    for (uint32_t i = threadIdx.x; i < (uint32_t)list_fwd_hists(MSCC) * S2; i += blockDim.x) lsm[i] = 0;
    __syncthreads();
    const uint32_t rl1 = P.read_len - 1;
    bool bad_order = false, bad_range = false;
    const uint32_t s_pair = smem_addr(h_pair), s_fp = smem_addr(h_fp), s_fm = smem_addr(h_fm);
    const uint32_t s_dump = s_pair + 4u * DUMP;         // (the pair histogram's own dump bin is never flushed)
    (void)s_fp; (void)s_fm; (void)s_dump;

    list_segments(P, warp, n_warps,
        [&](uint32_t, const ListJob& J, uint32_t w0, uint32_t w1) {
            const uint32_t* __restrict__ fwd = J.fwd;
            const uint32_t* __restrict__ rev = J.rev;
            const uint32_t* __restrict__ begins = J.begins;
            const uint32_t* __restrict__ ends = J.ends;
            const uint32_t nR = J.n_rev, nI = J.n_iv;
            const bool job_ncc = NCC && (J.flags & EVF_NCC);
            const bool job_mscc = MSCC && (J.flags & EVF_MSCC) && nI;
            const uint32_t r_last = nR ? __ldg(rev + nR - 1) : 0;       // (windows that end below it never meet the list's end)
            (void)r_last; (void)job_ncc;
            const uint32_t lo32 = (uint32_t)J.lo, hi32 = (uint32_t)J.hi;         // 0 <= lo <= hi <= nbits < 2^32
            const unsigned long long nbits = J.nbits;
            // max_shift as a value that has to live in a register: as a kernel parameter it was re-read
            // from the constant bank for every pair (LDCU + IMAD.U32); nbits < 2^32, so this adds 0
            const uint32_t Dr = D + (uint32_t)(nbits >> 32);
            PMB_CHECK(w0 < w1 && w1 <= J.n_fwd, P.flags);
            const uint32_t p_first = __ldg(fwd + w0);
            uint32_t JR = warp_lower_bound(rev, nR, p_first, lane);                          // reverse reads below p
            uint32_t CI = job_mscc ? warp_lower_bound(begins, nI, p_first + rl1, lane) : 0;  // runs beginning below p+rl-1
            uint32_t tot = 0;
            for (uint32_t base = w0; base < w1; base += 32) {
                const uint32_t i = base + lane;
                const uint32_t n_act = min(32u, w1 - base);
                const bool act = lane < (int)n_act;
                const uint32_t p = __ldg(fwd + (act ? i : w1 - 1));
                uint32_t pprev = __shfl_up_sync(0xFFFFFFFFu, p, 1);
                if (lane == 0) pprev = base ? __ldg(fwd + base - 1) : LIST_NONE;
                bad_order |= act && pprev != LIST_NONE && p < pprev;
                bad_range |= act && p >= nbits;
                const bool valid = act && p != pprev && p >= lo32 && p < hi32;       // (hi <= nbits < 2^32)
                tot += valid;
                const uint32_t p_max = __shfl_sync(0xFFFFFFFFu, p, n_act - 1);
                // this read's first reverse read: q >= p; the mappable runs: c = last run that begins
                // below x_top = p + rl - 1 (both searches' probes are issued before either is resolved)
                const uint32_t x_top = p + rl1;
                const uint32_t off_r = cursor_window(rev, nR, JR, p);
                const uint32_t off_i = job_mscc ? cursor_window(begins, nI, CI, x_top) : 0u;
                uint32_t j = cursor_resolve(rev, nR, JR, p, p_max, off_r);
                PMB_CHECK(j <= nR, P.flags);
                JR = __shfl_sync(0xFFFFFFFFu, j, n_act - 1);
                int c = -1;
                bool mp = false;
                if (job_mscc) {
                    const uint32_t c1 = cursor_resolve(begins, nI, CI, x_top, p_max + rl1, off_i);
                    PMB_CHECK(c1 <= nI, P.flags);
                    CI = __shfl_sync(0xFFFFFFFFu, c1, n_act - 1);
                    c = (int)c1 - 1;
                    int cp = c;
                    while (cp >= 0 && __ldg(begins + cp) >= p) cp--;
                    mp = valid && cp >= 0 && p <= __ldg(ends + cp);                          // M[p]
                }
                // every window of the group ends before the list does (whatever the input: no read of
                // the loop below can pass rev[n_rev - 1]) -> no bounds checks in the pair loop
                const bool fast = __all_sync(0xFFFFFFFFu, !valid || (j < nR && r_last > p && r_last - p > D));
                if (!valid) continue;
                // the run in the window that the next shifts can meet: x = x_top - k in (b, e]  <=>  k - k1 <= span
                // (no run, or an empty one: k1 = NONE, span = 0 -- no k passes); k2w: the last shift of
                // the run while runs are left below it, else NONE (nothing more to walk to)
                uint32_t k1 = LIST_NONE, span = 0, k2w = LIST_NONE;
                auto load_run = [&]() {
                    PMB_CHECK(c >= 0 && (uint32_t)c < nI, P.flags);
                    const uint32_t b = __ldg(begins + c), e = __ldg(ends + c);
                    const uint32_t k2 = x_top - b - 1;
                    k1 = x_top - min(e, x_top);
                    span = k2 - k1;
                    k2w = c > 0 ? k2 : LIST_NONE;
                    if (e > b) {                        // the run enters the window at k1 and leaves after k2
                        smem_inc(s_fp + 4u * min(k1, DUMP));
                        smem_inc(s_fm + 4u * min(k2 + 1, DUMP));
                    } else {
                        k1 = LIST_NONE; span = 0;
                    }
                };
                if (MSCC && mp && c >= 0) load_run(); else mp = false;
                // pairs (p, q), p <= q <= p + D.  Of equal reverse reads the LAST is counted (a bit is
                // set once, mscc.pyx:416-418): the test is against the entry loaded for the next round.
                // A pair that does not count goes to the dump bin.  `checked`: the list may end inside
                // this group's windows.
                auto pair_loop = [&](auto checked) {
                    bool more = !checked.value || j < nR;
                    uint32_t q = more ? __ldg(rev + j) : 0u;
                    uint32_t k = q - p;                 // (unsigned: an unsorted list cannot index outside the bins)
#pragma unroll 2
                    while (more && k <= Dr) {
                        ++j;
                        const bool last = checked.value && j >= nR;
                        const uint32_t qn = last ? ~q : __ldg(rev + j);
                        const uint32_t k4 = 4u * k;
                        const bool once = q != qn;
                        PMB_CHECK(k <= D && j <= nR, P.flags);
                        if (MSCC) {
                            while (k > k2w) { c--; load_run(); }
                            smem_add(s_pair + k4, once ? (k - k1 <= span ? 0x10001u : 1u) : 0u);
                        } else {
                            smem_inc(once ? s_pair + k4 : s_dump);
                        }
                        q = qn;
                        k = q - p;
                        more = !last;
                    }
                };
                if (fast) pair_loop(std::false_type{}); else pair_loop(std::true_type{});
                if (MSCC && mp)
                    while (k2w < D) { c--; load_run(); }              // the window's remaining runs
            }
            for (int o = 16; o; o >>= 1) tot += __shfl_xor_sync(0xFFFFFFFFu, tot, o);
            if (lane == 0 && tot) atomicAdd(J.acc + P.acc_off + (size_t)LA_N * P.stride, (unsigned long long)tot);   // bit_array_num_bits_set
        },
        [&](uint32_t, const ListJob& J) {
            unsigned long long* acc = J.acc + P.acc_off;
            const bool keep_ncc = NCC && (J.flags & EVF_NCC);     // (the pair loop counts for every job)
            if (MSCC) {
                for (uint32_t i = threadIdx.x; i <= D; i += blockDim.x) {
                    const uint32_t v = h_pair[i];
                    if (!v) continue;
                    if (keep_ncc) atomicAdd(acc + (size_t)LA_NCC * P.stride + i, (unsigned long long)(v & 0xFFFFu));
                    if (v >> 16) atomicAdd(acc + (size_t)LA_MSCC * P.stride + i, (unsigned long long)(v >> 16));
                    h_pair[i] = 0;
                }
                flush_hist(h_fp, D + 1, acc + (size_t)LA_FP * P.stride);
                flush_hist(h_fm, D + 1, acc + (size_t)LA_FM * P.stride);
                if (threadIdx.x == 0) { h_fp[DUMP] = 0; h_fm[DUMP] = 0; }
            } else if (keep_ncc) {
                flush_hist(h_pair, D + 1, acc + (size_t)LA_NCC * P.stride);
            } else {
                for (uint32_t i = threadIdx.x; i <= D; i += blockDim.x) h_pair[i] = 0;
            }
        }, MSCC ? (unsigned long long)min(max(P.pair_cap, 1u), LIST_PAIR_CAP) : ~0ull);
    if (bad_range) atomicExch(P.flags, 1);
    if (bad_order) atomicExch(P.flags + 1, 1);
}

// ---------------------------------------------------------------- reverse reads: d(mrev), strand total
template <bool MSCC>
__global__ void __launch_bounds__(LIST_THREADS, LIST_MIN_BLOCKS) list_rev_kernel(const ListParams P)
{
    extern __shared__ uint32_t lsm[];
    const uint32_t D = P.max_shift, S2 = D + 2;
    uint32_t* h_rp = lsm;
    uint32_t* h_rm = lsm + S2;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, n_warps = blockDim.x >> 5;
    if (MSCC) {
        for (uint32_t i = threadIdx.x; i < 2 * S2; i += blockDim.x) lsm[i] = 0;
        __syncthreads();
    }
    const uint32_t rl1 = P.read_len - 1;
    bool bad_order = false, bad_range = false;
    const uint32_t s_rp = smem_addr(h_rp), s_rm = smem_addr(h_rm);
    const uint32_t s_dump = s_rp + 4u * (D + 1);        // (bin D+1 of h_rp is never flushed)
    (void)s_rp; (void)s_rm; (void)s_dump;

    list_segments(P, warp, n_warps,
        [&](uint32_t, const ListJob& J, uint32_t w0, uint32_t w1) {
            const uint32_t* __restrict__ rev = J.rev;
            const uint32_t* __restrict__ begins = J.begins;
            const uint32_t* __restrict__ ends = J.ends;
            const uint32_t nI = J.n_iv;
            const bool job_mscc = MSCC && (J.flags & EVF_MSCC) && nI;
            const uint32_t lo32 = (uint32_t)J.lo, hi32 = (uint32_t)J.hi;         // 0 <= lo <= hi <= nbits < 2^32
            const unsigned long long nbits = J.nbits;
            PMB_CHECK(w0 < w1 && w1 <= J.n_rev, P.flags);
            uint32_t CI = job_mscc ? warp_lower_bound(begins, nI, __ldg(rev + w0) + rl1, lane) : 0;
            uint32_t tot = 0;
            for (uint32_t base = w0; base < w1; base += 32) {
                const uint32_t n_act = min(32u, w1 - base);
                const bool act = lane < (int)n_act;
                const uint32_t q = __ldg(rev + (act ? base + lane : w1 - 1));
                uint32_t qprev = __shfl_up_sync(0xFFFFFFFFu, q, 1);
                if (lane == 0) qprev = base ? __ldg(rev + base - 1) : LIST_NONE;
                bad_order |= act && qprev != LIST_NONE && q < qprev;
                bad_range |= act && q >= nbits;
                const bool valid = act && q != qprev && q < nbits;
                tot += valid && q >= lo32 && q < hi32;
                if (!job_mscc) continue;
                const uint32_t t = q + rl1;
                const uint32_t q_max = __shfl_sync(0xFFFFFFFFu, q, n_act - 1);
                const uint32_t c1 = cursor_lower_bound(begins, nI, CI, t, q_max + rl1);
                PMB_CHECK(c1 <= nI, P.flags);
                CI = __shfl_sync(0xFFFFFFFFu, c1, n_act - 1);
                // shifts whose j = q - k lies in the counted range
                const uint32_t ka = q >= hi32 ? min(q - hi32 + 1u, D + 1u) : 0u;
                const uint32_t kb = min(q - lo32, D);                  // (for q >= lo, checked below)
                int cB = (int)c1 - 1;                   // last run that begins below t = q + rl - 1
                int cA = cB;                            // ... below q
                while (cA >= 0 && __ldg(begins + cA) >= q) cA--;
                if (!valid || q < lo32 || ka > kb || cA < 0) continue;
                // M[q-k]: run (b, e] <=> k in [a1, a2];  M[t-2k]: run (b, e] <=> k in [b1, b2]
                uint32_t a1, a2, b1, b2;
                auto load_a = [&]() {
                    PMB_CHECK(cA >= 0 && (uint32_t)cA < nI, P.flags);
                    const uint32_t b = __ldg(begins + cA), e = __ldg(ends + cA);
                    a1 = q - min(e, q); a2 = q - b - 1;
                };
                auto load_b = [&]() {
                    PMB_CHECK(cB >= 0 && (uint32_t)cB < nI, P.flags);
                    const uint32_t b = __ldg(begins + cB), e = __ldg(ends + cB);
                    b1 = (t - min(e, t) + 1) >> 1; b2 = (t - b - 1) >> 1;
                };
                load_a();
                load_b();
                for (;;) {
                    if (a1 > kb || b1 > kb) break;
                    const uint32_t s = max(max(a1, b1), ka), e = min(min(a2, b2), kb);
                    {                                   // both mappable for k in [s, e]; else: the dump bin
                        const bool both = s <= e;
                        PMB_CHECK(!both || e <= D, P.flags);
                        smem_inc(both ? s_rp + 4u * s : s_dump);
                        smem_inc(both ? s_rm + 4u * e + 4u : s_dump);
                    }
                    if (min(a2, b2) >= kb) break;
                    if (a2 <= b2) { if (cA == 0) break; cA--; load_a(); }
                    else { if (cB == 0) break; cB--; load_b(); }
                }
            }
            for (int o = 16; o; o >>= 1) tot += __shfl_xor_sync(0xFFFFFFFFu, tot, o);
            if (lane == 0 && tot) atomicAdd(J.acc + P.acc_off + (size_t)LA_N * P.stride + 1, (unsigned long long)tot);
        },
        [&](uint32_t, const ListJob& J) {
            if (MSCC) {
                flush_hist(h_rp, D + 1, J.acc + P.acc_off + (size_t)LA_RP * P.stride);
                flush_hist(h_rm, D + 1, J.acc + P.acc_off + (size_t)LA_RM * P.stride);
                if (threadIdx.x == 0) h_rm[D + 1] = 0;
            }
        });
    if (bad_range) atomicExch(P.flags, 1);
    if (bad_order) atomicExch(P.flags + 1, 1);
}

// ---------------------------------------------------------------- mappable runs: edge products, direct lags
// (entries = the intervals of the jobs that OWN a track; followers and NCC-only jobs have none)
__global__ void __launch_bounds__(LIST_THREADS, LIST_MIN_BLOCKS) list_edge_kernel(const ListParams P)
{
    extern __shared__ uint32_t lsm[];
    const uint32_t D = P.max_shift, S2 = D + 2;
    uint32_t* h_plus = lsm;
    uint32_t* h_minus = lsm + S2;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, n_warps = blockDim.x >> 5;
    for (uint32_t i = threadIdx.x; i < 2 * S2; i += blockDim.x) lsm[i] = 0;
    __syncthreads();
    bool bad_order = false, bad_range = false;
    const uint32_t s_plus = smem_addr(h_plus), s_minus = smem_addr(h_minus);
    const uint32_t s_dump = s_plus + 4u * (D + 1);      // (bin D+1 of h_plus is never flushed)

    list_segments(P, warp, n_warps,
        [&](uint32_t, const ListJob& J, uint32_t w0, uint32_t w1) {
            const uint32_t* __restrict__ begins = J.begins;
            const uint32_t* __restrict__ ends = J.ends;
            const uint32_t n = J.n_iv;
            const bool autocorr = (J.flags & EVF_AUTOCORR) != 0;
            const long long lo = J.lo, hi = J.hi;
            const long long off = autocorr ? 0 : (long long)P.read_len - 1 - (long long)D;   // window start relative to an edge
            const uint32_t L0 = autocorr ? 0u : P.read_len - 1u;         // (read_len >= 2 on this path)
            const uint32_t L1 = autocorr ? 1u : P.read_len - 2u;
            const uint32_t Lmax = L0 > L1 ? L0 : L1;
            unsigned long long d0 = 0, d1 = 0;
            for (uint32_t base = w0; base < w1; base += 32) {
                const uint32_t i = base + lane;
                if (i >= w1) continue;
                PMB_CHECK(i < n, P.flags);
                const long long b = __ldg(begins + i), e = __ldg(ends + i);
                bad_order |= b > e || (i + 1 < n && e > (long long)__ldg(begins + i + 1));
                bad_range |= e > b && (unsigned long long)e >= J.nbits;
                const long long s = max(b + 1, lo), t = min(e + 1, hi);     // the counted part of the run: [s, t)
                if (s >= t) continue;
                // run edges of M inside the windows of this run's two edges: [s + off, s + off + D] and [t + off, ...]
                // From here on 32-bit arithmetic: the list path is not taken for a chromosome within
                // 2 (max_shift + read_len) of 2^32 bits (pmb_api.cu, list_ok), so the sums below fit, and a
                // distance taken modulo 2^32 is in the window iff it is <= D as an unsigned number (neither
                // a negative nor a huge true distance can wrap into [0, D]).
                const uint32_t w_hi = (uint32_t)(t + off + (long long)D);           // >= 1: off + D >= 0
                const uint32_t key = (uint32_t)max(s + off - 1, 0ll);
                // the first run whose end reaches the start edge's window: ends[Jx] >= s + off - 1.  It lies a
                // run or two from this one -- walked to, one entry at a time (never more steps than the loop
                // below then makes)
                uint32_t Jx = i;
                while (Jx > 0 && __ldg(ends + Jx - 1) >= key) --Jx;
                if (Jx == i)
                    while (Jx < n && __ldg(ends + Jx) < key) ++Jx;
                const uint32_t base0 = (uint32_t)s + (uint32_t)off, base1 = (uint32_t)t + (uint32_t)off;
                for (; Jx < n; ++Jx) {
                    const uint32_t bj = __ldg(begins + Jx), ej = __ldg(ends + Jx);
                    if (bj + 1u > w_hi) break;
                    if (ej <= bj) continue;
                    const uint32_t ds0 = bj + 1u - base0, de0 = ej + 1u - base0;     // against the start edge
                    const uint32_t ds1 = bj + 1u - base1, de1 = ej + 1u - base1;     // against the end edge
                    smem_inc(ds0 <= D ? s_plus + 4u * ds0 : s_dump);                // start . start
                    smem_inc(de0 <= D ? s_minus + 4u * de0 : s_dump);               // start . end
                    smem_inc(ds1 <= D ? s_minus + 4u * ds1 : s_dump);               // end . start
                    smem_inc(de1 <= D ? s_plus + 4u * de1 : s_dump);                // end . end
                }
                // the two directly counted lags: overlap of this run with the runs at and above it, moved down
                const uint32_t A_lo = (uint32_t)s, A_hi = (uint32_t)t - 1u;
                for (uint32_t kk = i; kk < n; ++kk) {
                    const uint32_t B_lo = __ldg(begins + kk) + 1u, B_hi = __ldg(ends + kk);
                    if (B_lo > A_hi + Lmax) break;
                    if (B_hi < B_lo) continue;
                    if (B_hi >= L0) {
                        const uint32_t top = min(A_hi, B_hi - L0), bot = max(A_lo, B_lo > L0 ? B_lo - L0 : 0u);
                        if (top >= bot) d0 += top - bot + 1u;
                    }
                    if (B_hi >= L1) {
                        const uint32_t top = min(A_hi, B_hi - L1), bot = max(A_lo, B_lo > L1 ? B_lo - L1 : 0u);
                        if (top >= bot) d1 += top - bot + 1u;
                    }
                }
            }
            for (int o = 16; o; o >>= 1) {
                d0 += __shfl_xor_sync(0xFFFFFFFFu, d0, o);
                d1 += __shfl_xor_sync(0xFFFFFFFFu, d1, o);
            }
            if (lane == 0) {
                unsigned long long* lags = J.aux + P.acc_off + 2 * (size_t)P.stride;
                if (d0) atomicAdd(lags, d0);
                if (d1) atomicAdd(lags + 1, d1);
            }
        },
        [&](uint32_t, const ListJob& J) {
            flush_hist(h_plus, D + 1, J.aux + P.acc_off);
            flush_hist(h_minus, D + 1, J.aux + P.acc_off + P.stride);
        });
    if (bad_range) atomicExch(P.flags, 1);
    if (bad_order) atomicExch(P.flags + 1, 1);
}

// ---------------------------------------------------------------- finalize: accumulators -> outputs
// Inclusive prefix sum of a[0..n) in place by one CTA of 256 threads: every thread owns a
// contiguous chunk, the 256 chunk sums are scanned by warp shuffles (a serial pass of one thread
// over them cost 4 us per scan: the whole kernel was latency, 27 us whatever the batch).
__device__ __forceinline__ long long shfl_up_ll(long long v, int d)
{
    const int lo = __shfl_up_sync(0xFFFFFFFFu, (int)(unsigned long long)v, d);
    const int hi = __shfl_up_sync(0xFFFFFFFFu, (int)((unsigned long long)v >> 32), d);
    return (long long)(((unsigned long long)(unsigned)hi << 32) | (unsigned)lo);
}
__device__ __forceinline__ void block_scan_fast(long long* a, uint32_t n, long long* warp_sums /* [8] shared */)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t chunk = (n + blockDim.x - 1) / blockDim.x;
    const uint32_t lo = min(n, threadIdx.x * chunk), hi = min(n, lo + chunk);
    long long sum = 0;
    for (uint32_t i = lo; i < hi; ++i) sum += a[i];
    long long inc = sum;                                  // inclusive scan over the warp
    for (int d = 1; d < 32; d <<= 1) {
        const long long t = shfl_up_ll(inc, d);
        if (lane >= d) inc += t;
    }
    if (lane == 31) warp_sums[warp] = inc;
    __syncthreads();
    long long run = inc - sum;                            // exclusive prefix inside the warp
    for (int w = 0; w < warp; ++w) run += warp_sums[w];
    for (uint32_t i = lo; i < hi; ++i) { run += a[i]; a[i] = run; }
    __syncthreads();
}

// grid (n_jobs, 4): part 0 ncc, mscc, strand totals (copies) | 1 mfwd | 2 mrev (prefix sums of their
// difference histograms) | 3 dmp (double prefix sum of the run-edge products, anchored by the two
// direct lags).  Outputs are written with plain stores; then the part clears the accumulators the
// NEXT execute adds into (see ListJob).  A job outside this rank's position range has its vectors
// written as zero (nothing adds to it, but an in-place all-reduce overwrites them).
__global__ void __launch_bounds__(256) list_finalize_kernel(const ListParams P)
{
    __shared__ long long warp_sums[8];
    const ListJob& J = P.jobs[blockIdx.x];
    const uint32_t part = blockIdx.y, stride = P.stride, D = P.max_shift, S = D + 1;
    if (blockIdx.x == 0 && part == 0 && threadIdx.x == 0) {
        if (P.status_tail) {            // this rank's status words, summed with the counters by pmb_allreduce
            P.status_tail[0] = P.flags[0] != 0 ? 1ull : 0ull;
            P.status_tail[1] = P.flags[1] != 0 ? 1ull : 0ull;
        }
        if (P.flags_other) { P.flags_other[0] = 0; P.flags_other[1] = 0; }
    }
    if (J.lo >= J.hi) {
        // A chromosome outside this rank's position range: nothing was added for it, its vectors are
        // zero -- and have to be WRITTEN as zero: an in-place all-reduce (pmb_allreduce) leaves the
        // other ranks' sums in them, which the next execute + reduce of the batch would add again.
        if (part == 0) {
            for (uint32_t k = threadIdx.x; k < S; k += blockDim.x) {
                J.counters[(size_t)C_NCC * stride + k] = 0;
                J.counters[(size_t)C_MSCC * stride + k] = 0;
            }
            if (threadIdx.x < 2) J.totals[threadIdx.x] = 0;
        } else {
            const int which = part == 1 ? C_MFWD : part == 2 ? C_MREV : C_DMP;
            for (uint32_t k = threadIdx.x; k < S; k += blockDim.x) J.counters[(size_t)which * stride + k] = 0;
        }
        return;
    }
    const bool autocorr = (J.flags & EVF_AUTOCORR) != 0;
    const bool mscc = (J.flags & EVF_MSCC) != 0;
    const bool clear = P.acc_other != P.acc_off;
    const long long* acc = (const long long*)(J.acc + P.acc_off);
    unsigned long long* acc_next = J.acc + P.acc_other;
    auto zero = [&](unsigned long long* p, uint32_t n) {
        if (clear) for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) p[i] = 0;
    };
    if (part == 0) {
        if (!autocorr) {
            if (J.flags & EVF_NCC)
                for (uint32_t k = threadIdx.x; k < S; k += blockDim.x) J.counters[(size_t)C_NCC * stride + k] = acc[(size_t)LA_NCC * stride + k];
            if (mscc)
                for (uint32_t k = threadIdx.x; k < S; k += blockDim.x) J.counters[(size_t)C_MSCC * stride + k] = acc[(size_t)LA_MSCC * stride + k];
            if (threadIdx.x < 2) J.totals[threadIdx.x] = acc[(size_t)LA_N * stride + threadIdx.x];
        }
        zero(acc_next + (size_t)LA_NCC * stride, S);
        zero(acc_next + (size_t)LA_MSCC * stride, S);
        zero(acc_next + (size_t)LA_N * stride, 2);
        return;
    }
    if (part == 1 || part == 2) {
        const int lp = part == 1 ? LA_FP : LA_RP;
        if (mscc && !autocorr) {
            long long* out = (long long*)(J.counters + (size_t)(part == 1 ? C_MFWD : C_MREV) * stride);
            const long long* plus = acc + (size_t)lp * stride;
            const long long* minus = plus + stride;
            for (uint32_t k = threadIdx.x; k < S; k += blockDim.x) out[k] = plus[k] - minus[k];
            __syncthreads();
            block_scan_fast(out, S, warp_sums);
        }
        zero(acc_next + (size_t)lp * stride, S);
        zero(acc_next + (size_t)(lp + 1) * stride, S);
        return;
    }
    // part 3: dmp
    if (mscc || autocorr) {
        long long* out = (long long*)(J.counters + (size_t)C_DMP * stride);
        const long long* plus = (const long long*)(J.aux + P.acc_off);
        const long long* minus = plus + stride;
        const long long a0 = plus[2 * (size_t)stride], a1 = plus[2 * (size_t)stride + 1];
        if (D == 0) {
            if (threadIdx.x == 0) out[0] = a0;
        } else {
            // out[0] = A0, out[1] = A1 - A0, out[k+1] = -C[b(k)]; two scans: first differences, then dmp
            for (uint32_t k = threadIdx.x; k <= D; k += blockDim.x) {
                long long v;
                if (k == 0) v = a0;
                else if (k == 1) v = a1 - a0;
                else { const uint32_t b = autocorr ? k - 1 : D - (k - 1); v = -(plus[b] - minus[b]); }
                out[k] = v;
            }
            __syncthreads();
            block_scan_fast(out + 1, D, warp_sums);
            block_scan_fast(out, D + 1, warp_sums);
        }
    }
    if (!(J.flags & EVF_SHARED)) {      // the owner clears the run-edge accumulators its followers read, too
        unsigned long long* aux_next = J.aux + P.acc_other;
        zero(aux_next, S);
        zero(aux_next + stride, S);
        zero(aux_next + 2 * (size_t)stride, 2);
    }
}

}  // namespace pmb
