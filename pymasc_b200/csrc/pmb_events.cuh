// pmb_events.cuh -- the event-window path (sm_100a): the same per-shift counts as the
// shift-window kernel, computed from the SET BITS of the sparse arrays instead of from every
// bit position.
//
// The reference's loop (PyMaSC/core/bitarray/mscc.pyx:288-317) and the dense shift-window
// kernel (pmb_window.cuh) cost  L x (max_shift+1)  whatever the data.  But F and R hold one
// bit per read (0.3 % of the positions at 20 M reads on hg38) and M is a union of runs, so
// every count is a sum over few "events", each contributing one (max_shift+1)-bit WINDOW of a
// bit array (bit b of the window <-> shift k = b):
//
//   forward read at p:   ncc  += R[p .. p+D]                                   (:314)
//     and, if M[p]:      mfwd += M[p+rl-1-k], k = 0..D   (descending window)   (:303)
//                        mscc += R[p+k] & M[p+rl-1-k]                          (:305)
//   reverse read at q:   mrev += M[q-k] & M[q+rl-1-2k]   (j = q-k in range)    (:304)
//                             =  M[q-k] & {Er|Or}[N/2-1-((q+rl-1)>>1) .. +D]
//   mappable-run edges:  popc(D_k) = A[rl-1-k],  A[s] = sum_j Mc[j] M[j+s]     (:291-298)
//                        2A[s] - A[s-1] - A[s+1] = sum_j dMc[j] dM[j+s]  (dX = starts - ends),
//                        so every run edge of the (range-clipped) M adds the window of run
//                        starts / ends of M (formed on the fly from two funnel shifts of M);
//                        two directly counted lags anchor the recurrence (finalize_dmp_kernel).
//                        Exact integer arithmetic.
//
// Descending windows are read from M itself (per-lane descending words + brev).  Er/Or are the
// even/odd bits of M bit-reversed (the stride-2 descending window becomes an ascending one);
// derive_kernel writes them in one streaming pass over M.
//
// Mapping.  A CTA stages a tile of the event source and the window regions of that tile in
// shared memory by TMA (see "Tile staging" below); a warp takes 32 source words at a time and
// handles their events one after the other: lane <-> window column (32 shifts), so one event is two
// conflict-free LDS + one funnel shift per lane and array.  The per-shift
// counts are kept BIT-SLICED in registers: a carry-save binary counter (acc/pend word per
// level, ~1 CSA = 2 LOP3 per event and counter, independent of the window's density); a
// carry out of the top level and the final residue go to a per-block shared-memory
// histogram, which is flushed to the 64-bit global counters when the block changes
// chromosome or ends.  No data-dependent approximation: counts are bit-exact.
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>

#include "pmb_window.cuh"   // xor3 / maj3, counter slots

namespace pmb {

enum { EV_FWD = 0, EV_REV = 1, EV_EDGE = 2 };
enum { EVF_NCC = 1u, EVF_MSCC = 2u, EVF_AUTOCORR = 4u,
       EVF_SHARED = 8u };   // M, Er/Or, aux and the direct lags belong to an earlier job (same track)

struct EventJob {
    const uint32_t* F;    // array word 0 of each array (halo included)
    const uint32_t* R;
    const uint32_t* M;
    const uint32_t* Er;   // even / odd bits of M, bit-reversed (half arrays)
    const uint32_t* Or;
    unsigned long long* counters;   // [5][stride]
    unsigned long long* aux;        // [2][stride]: edge products, + then - (window-bit indexed)
    unsigned long long* adirect;    // [2]: the two directly counted lags
    unsigned long long* totals;     // [2]: popc(F), popc(R) over the counted range
    long long a_lo, a_hi;           // counted range in array bits
    uint32_t array_words;
    uint32_t tile_begin, n_tiles;
    int32_t edge_off;               // window start relative to an edge: rl-1-D, or 0 (autocorr)
    uint32_t flags;
    uint32_t gate_index;            // H2D group of the job: which "unsorted" flag covers it
    // gather build: the direct lags counted from the intervals (valid unless the flag is set)
    const unsigned long long* agather;
};

struct EventParams {
    const EventJob* jobs;
    uint32_t n_jobs;
    uint32_t tile_lo;       // this launch covers the tiles [tile_lo, tile_hi) of the batch
    uint32_t tile_hi;       //   (everything, or one H2D group of jobs)
    uint32_t tile_words;    // multiple of 32 * warps per block
    uint32_t col0;          // first window column of this launch
    uint32_t n_cols;        // <= 32
    uint32_t max_shift;
    uint32_t read_len;
    uint32_t stride;
    const int* gate;        // EV_EDGE: run only if *gate is set (fallback of edge_list_kernel); may be null
    const uint32_t* mem_lo; // the bit-array allocation (pads included): every TMA source range lies inside
    const uint32_t* mem_hi; //   (verified in a -DPMB_CHECKED build)
    int* chk;               // the batch's error flag
};

#ifndef EV_SWITCH_ADD
#define EV_SWITCH_ADD 1
#endif
constexpr int EV_LEVELS = 6;          // carry-save levels: a top-level carry every 64 events
constexpr int EV_HIST_STRIDE = 33;    // padded: lane l, bit b -> bank (l + b) % 32

// NC bit-sliced counters sharing one event count (all warp-uniform control flow).
template <int NC>
struct Slices {
    uint32_t acc[NC][EV_LEVELS];
    uint32_t pend[NC][EV_LEVELS];
    uint32_t m;

    __device__ __forceinline__ void reset()
    {
#pragma unroll
        for (int n = 0; n < NC; ++n)
#pragma unroll
            for (int i = 0; i < EV_LEVELS; ++i) { acc[n][i] = 0; pend[n][i] = 0; }
        m = 0;
    }
    static __device__ __forceinline__ void sink(uint32_t* hist, int lane, uint32_t c, uint32_t weight)
    {
        while (c) {
            const int b = __ffs(c) - 1;
            c &= c - 1;
            atomicAdd(hist + lane * EV_HIST_STRIDE + b, weight);
        }
    }
    // Event number m with TZ trailing zero bits: levels 0..TZ-1 each fold (acc, pend, carry)
    // with one CSA, level TZ parks the carry (or, past the top level, it goes to the histogram).
    template <int TZ>
    __device__ __forceinline__ void add_case(const uint32_t (&x)[NC], uint32_t* hist, int lane)
    {
        uint32_t c[NC];
#pragma unroll
        for (int n = 0; n < NC; ++n) c[n] = x[n];
#pragma unroll
        for (int i = 0; i < TZ; ++i) {
#pragma unroll
            for (int n = 0; n < NC; ++n) {
                const uint32_t t = maj3(acc[n][i], pend[n][i], c[n]);
                acc[n][i] = xor3(acc[n][i], pend[n][i], c[n]);
                c[n] = t;
            }
        }
        if (TZ < EV_LEVELS) {
#pragma unroll
            for (int n = 0; n < NC; ++n) pend[n][TZ < EV_LEVELS ? TZ : 0] = c[n];
        } else {
#pragma unroll
            for (int n = 0; n < NC; ++n) sink(hist + n * 32 * EV_HIST_STRIDE, lane, c[n], 1u << EV_LEVELS);
        }
    }
    // add one window word per counter; m is warp-uniform, so the switch is a uniform branch
    // (a predicated cascade over all levels costs ~6x the instructions)
    __device__ __forceinline__ void add(const uint32_t (&x)[NC], uint32_t* hist, int lane)
    {
        ++m;
        static_assert(EV_LEVELS == 6, "switch below lists the cases of 6 levels");
        const int tz = __ffs(m | (1u << EV_LEVELS)) - 1;
#if EV_SWITCH_ADD == 2
        if (tz == 0) add_case<0>(x, hist, lane);
        else if (tz == 1) add_case<1>(x, hist, lane);
        else if (tz == 2) add_case<2>(x, hist, lane);
        else if (tz == 3) add_case<3>(x, hist, lane);
        else if (tz == 4) add_case<4>(x, hist, lane);
        else if (tz == 5) add_case<5>(x, hist, lane);
        else add_case<6>(x, hist, lane);
        return;
#endif
        switch (tz) {
            case 0: add_case<0>(x, hist, lane); break;
            case 1: add_case<1>(x, hist, lane); break;
            case 2: add_case<2>(x, hist, lane); break;
            case 3: add_case<3>(x, hist, lane); break;
            case 4: add_case<4>(x, hist, lane); break;
            case 5: add_case<5>(x, hist, lane); break;
            default: add_case<6>(x, hist, lane); break;
        }
    }
    // The same, with the window words produced by `compute(x)`.  The trailing zeros of m are
    // tested bit by bit (half of all events stop at the first test and just park x in pend[0];
    // the jump-table dispatch of a switch costs ~18 instructions for every event).
    template <class F>
    __device__ __forceinline__ void add_with(uint32_t* hist, int lane, F&& compute)
    {
        static_assert(EV_LEVELS == 6, "the chain below lists the cases of 6 levels");
        ++m;
        if (m & 1u) {           // the words are produced straight into the level-0 slots
            uint32_t x[NC];
            compute(x);
#pragma unroll
            for (int n = 0; n < NC; ++n) pend[n][0] = x[n];
            return;
        }
        // (compute is instantiated a second time here on purpose: with one shared copy the
        // compiler computes x into pend[.][0]'s registers and pays 9 moves per event to save,
        // clear and restore them around the carry-save cases)
        uint32_t x[NC];
        compute(x);
        if (m & 2u) add_case<1>(x, hist, lane);
        else if (m & 4u) add_case<2>(x, hist, lane);
        else if (m & 8u) add_case<3>(x, hist, lane);
        else if (m & 16u) add_case<4>(x, hist, lane);
        else if (m & 32u) add_case<5>(x, hist, lane);
        else add_case<6>(x, hist, lane);
    }
    __device__ __forceinline__ void flush(uint32_t* hist, int lane)
    {
#pragma unroll
        for (int n = 0; n < NC; ++n) {
#pragma unroll
            for (int i = 0; i < EV_LEVELS; ++i) {
                sink(hist + n * 32 * EV_HIST_STRIDE, lane, acc[n][i], 1u << i);
                if ((m >> i) & 1u) sink(hist + n * 32 * EV_HIST_STRIDE, lane, pend[n][i], 1u << i);
            }
        }
        reset();
    }
};

__device__ __forceinline__ uint32_t range_mask(long long bit0, long long lo, long long hi)
{
    if (bit0 + 32 <= lo || bit0 >= hi) return 0u;
    uint32_t mask = 0xFFFFFFFFu;
    if (bit0 < lo) mask &= 0xFFFFFFFFu << (lo - bit0);
    if (bit0 + 32 > hi) mask &= 0xFFFFFFFFu >> (bit0 + 32 - hi);
    return mask;
}

// Tile staging.  All events of one 32-word scan group (1024 bit positions from array bit G)
// need windows that lie in a <= 66-word span of each array, whose start depends on G only:
//   ascending arrays  (R, Sb, Eb):  window start = B + e,          B = G + c0
//   reversed arrays   (Mr, Er, Or): window start = B + (emax - e), B = C - G - emax
// with e the event's offset inside the group.  So a TILE of EV_TILE source words needs one
// contiguous region of every array, and a CTA brings those regions in with 1-D TMA bulk copies
// (cp.async.bulk.shared::cluster.global, completing on an mbarrier), double buffered: the
// regions of the next tile are in flight while the warps walk the events of this one.  An
// event then costs two conflict-free LDS and one funnel shift per lane and array, and global
// latency is off the critical path.  Regions may overhang the job's array by a few words on
// either side (the allocation is padded; such words are never used by a valid event).
constexpr int EV_TILE = 128;                  // source words per (warp-private) tile = 4 scan groups
constexpr int EV_WARPS = 8;                   // warps per CTA
constexpr int EV_REGION = EV_TILE + 40;       // words of a full-rate window region
constexpr int EV_REGION_HALF = EV_TILE / 2 + 44;   // decimated arrays (Er, Or)
constexpr int EV_PAD_WORDS = 2048;            // padding of the bit-array allocation, both ends

template <int KIND>
struct EvTraits;
template <> struct EvTraits<EV_FWD> {     // counters ncc, mfwd, mscc; tiles F, M | regions R, M
    static constexpr int NC = 3;
    static constexpr int STAGE_WORDS = 2 * EV_TILE + 2 * EV_REGION;
};
template <> struct EvTraits<EV_REV> {     // counter mrev; tile R | regions M, Er, Or
    static constexpr int NC = 1;
    static constexpr int STAGE_WORDS = EV_TILE + EV_REGION + 2 * EV_REGION_HALF;
};
template <> struct EvTraits<EV_EDGE> {    // + and - edge products; tile M (4 words early) | region M
    static constexpr int NC = 2;
    static constexpr int STAGE_WORDS = (EV_TILE + 4) + EV_REGION;
};

template <int KIND>
__host__ __device__ constexpr size_t event_smem_bytes()
{
    return 16 * EV_WARPS + (size_t)EvTraits<KIND>::NC * 32 * EV_HIST_STRIDE * 4 +
           2 * (size_t)EV_WARPS * EvTraits<KIND>::STAGE_WORDS * 4;
}

__device__ __forceinline__ long long floor4(long long w) { return w & ~3ll; }

// window column `lane` at bit offset o of a staged span
__device__ __forceinline__ uint32_t span_word(const uint32_t* sp, uint32_t o, int lane)
{
    const uint32_t w = (o >> 5) + lane;
    return __funnelshift_r(sp[w], sp[w + 1], o & 31);
}

// Descending window: bit k of the result is A[top - k] for lane 0's top = o + 31, lane l 32 l lower.
// (window column `lane` of the bit-reversed array, read from the array itself: no reversed copy)
__device__ __forceinline__ uint32_t span_word_rev(const uint32_t* sp, uint32_t o, int lane)
{
    const uint32_t w = (o >> 5) - lane;
    return __brev(__funnelshift_r(sp[w], sp[w + 1], o & 31));
}

// KIND: which events; NCC / MSCC: which counters the batch asked for.
template <int KIND, bool NCC, bool MSCC>
__global__ void __launch_bounds__(32 * EV_WARPS) event_window_kernel(const EventParams P)
{
    constexpr int NC = EvTraits<KIND>::NC;
    constexpr int STAGE = EvTraits<KIND>::STAGE_WORDS;
    extern __shared__ __align__(16) unsigned char ev_smem[];
    uint64_t* full_all = reinterpret_cast<uint64_t*>(ev_smem);              // [EV_WARPS][2]
    uint32_t* hist = reinterpret_cast<uint32_t*>(ev_smem + 16 * EV_WARPS);  // [NC][32][33]
    uint32_t* stage_all = hist + NC * 32 * EV_HIST_STRIDE;                  // [EV_WARPS][2][STAGE]

    const int lane = threadIdx.x & 31;
    // read through a shuffle so the compiler KNOWS it is warp-uniform: the tile loops, the event
    // count and the carry-save dispatch then compile to uniform branches (BRA.U, no BSSY/BSYNC)
    const int warp = __shfl_sync(0xFFFFFFFFu, (int)(threadIdx.x >> 5), 0);

    if (KIND == EV_EDGE && P.gate && !*P.gate) return;
    const uint32_t n_launch_tiles = P.tile_hi - P.tile_lo;
    const uint32_t t_lo = P.tile_lo + (uint32_t)(((uint64_t)n_launch_tiles * blockIdx.x) / gridDim.x);
    const uint32_t t_hi = P.tile_lo + (uint32_t)(((uint64_t)n_launch_tiles * (blockIdx.x + 1)) / gridDim.x);
    if (t_lo >= t_hi) return;

    for (uint32_t i = threadIdx.x; i < NC * 32 * EV_HIST_STRIDE; i += blockDim.x) hist[i] = 0;
    // every warp runs its own two-stage TMA pipeline (no CTA barrier per tile: events per tile
    // vary, and a shared barrier left warps idle 3-5 issue slots per instruction)
    uint64_t* full = full_all + 2 * warp;
    uint32_t* stage0 = stage_all + (size_t)2 * warp * STAGE;
    if (lane == 0) {
        mbar_init(&full[0], 1);
        mbar_init(&full[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    // always 0, but not provably so for ptxas: adding it moves a warp-uniform value (and all the
    // arithmetic that follows from it) from the uniform datapath to the vector ALU
    // (read back from the freshly zeroed histogram: anything computed from %laneid is folded)
    const uint32_t vector_zero = hist[lane];

    Slices<NC> sl;
    sl.reset();
    uint32_t tot_f = 0, tot_r = 0;       // EV_FWD, first column chunk: strand totals of the job
    const uint32_t col_mask = lane < (int)P.n_cols ? 0xFFFFFFFFu : 0u;   // idle lanes when n_cols < 32
    const int D = (int)P.max_shift;
    const int rl = (int)P.read_len;
    const int hmax = (1022 + rl) >> 1;

    auto job_active = [&](const EventJob& J) -> bool {
        if (KIND != EV_EDGE && (J.flags & EVF_AUTOCORR)) return false;     // no reads in such a job
        if (KIND == EV_REV && !(J.flags & EVF_MSCC)) return false;
        if (KIND == EV_EDGE && !(J.flags & (EVF_MSCC | EVF_AUTOCORR))) return false;
        if (KIND == EV_EDGE && (J.flags & EVF_SHARED)) return false;       // the owner's scan covers it
        return true;
    };
    // Region origins (array word index, multiple of 4) of tile word offset g0 of job J.
    // a: first full-rate window region, b: decimated regions (REV only).
    auto region_a = [&](const EventJob& J, uint32_t g0) -> long long {
        // the M region of the tile: lowest word any lane touches for any event of the tile
        const long long G0 = (long long)g0 * 32;
        if (KIND == EV_FWD) return floor4(((G0 + rl - 32) >> 5) - P.col0 - 31);   // M[p+rl-1-k], descending
        if (KIND == EV_REV) return floor4(((G0 - 31) >> 5) - P.col0 - 31);        // M[q-k], descending
        return floor4(((G0 + J.edge_off - 1) >> 5) + P.col0);                     // dM[j+s], ascending
    };
    auto region_b = [&](const EventJob& J, uint32_t g0) -> long long {
        const long long N = (long long)J.array_words * 32;
        const long long G_last = ((long long)g0 + EV_TILE - 32) * 32;
        return floor4((((N >> 1) - 1 - (G_last >> 1) - hmax) >> 5) + P.col0);               // Er, Or
    };

    // (checked build: the source range of every bulk copy lies inside the bit-array allocation,
    // the destination inside this warp's stage)
    auto tma_load_1d = [&](uint32_t* dst, const uint32_t* src, uint32_t bytes, uint64_t* bar) {
        PMB_CHECK(src >= P.mem_lo && src + bytes / 4 <= P.mem_hi && (bytes & 15u) == 0 && ((uintptr_t)src & 15u) == 0, P.chk);
        PMB_CHECK(dst >= stage0 && dst + bytes / 4 <= stage0 + 2 * STAGE, P.chk);
        pmb::tma_load_1d(dst, src, bytes, bar);
    };
    // producer (thread 0): request the regions of one tile into one stage
    auto request = [&](uint32_t t, uint32_t job, int s) {
        const EventJob& J = P.jobs[job];
        const uint32_t g0 = (t - J.tile_begin) * EV_TILE;
        uint32_t* st = stage0 + (size_t)s * STAGE;
        fence_proxy_async();
        if (KIND == EV_FWD) {
            const bool mscc = MSCC && (J.flags & EVF_MSCC);
            mbar_expect_tx(&full[s], (EV_TILE + EV_REGION + (mscc ? EV_TILE + EV_REGION : 0)) * 4);
            tma_load_1d(st, J.F + g0, EV_TILE * 4, &full[s]);
            tma_load_1d(st + 2 * EV_TILE, J.R + g0 + P.col0, EV_REGION * 4, &full[s]);
            if (mscc) {
                tma_load_1d(st + EV_TILE, J.M + g0, EV_TILE * 4, &full[s]);
                tma_load_1d(st + 2 * EV_TILE + EV_REGION, J.M + region_a(J, g0), EV_REGION * 4, &full[s]);
            }
        } else if (KIND == EV_REV) {
            mbar_expect_tx(&full[s], (EV_TILE + EV_REGION + 2 * EV_REGION_HALF) * 4);
            tma_load_1d(st, J.R + g0, EV_TILE * 4, &full[s]);
            tma_load_1d(st + EV_TILE, J.M + region_a(J, g0), EV_REGION * 4, &full[s]);
            const long long rb = region_b(J, g0);
            tma_load_1d(st + EV_TILE + EV_REGION, J.Er + rb, EV_REGION_HALF * 4, &full[s]);
            tma_load_1d(st + EV_TILE + EV_REGION + EV_REGION_HALF, J.Or + rb, EV_REGION_HALF * 4, &full[s]);
        } else {
            mbar_expect_tx(&full[s], ((EV_TILE + 4) + EV_REGION) * 4);
            tma_load_1d(st, J.M + (long long)g0 - 4, (EV_TILE + 4) * 4, &full[s]);
            tma_load_1d(st + EV_TILE + 4, J.M + region_a(J, g0), EV_REGION * 4, &full[s]);
        }
    };
    auto flush_block = [&](uint32_t job) {
        sl.flush(hist, lane);
        __syncthreads();
        const EventJob& J = P.jobs[job];
        if (KIND == EV_FWD && P.col0 == 0) {      // bit_array_num_bits_set (bit_array.c:962-978)
            for (int o = 16; o; o >>= 1) {
                tot_f += __shfl_xor_sync(0xFFFFFFFFu, tot_f, o);
                tot_r += __shfl_xor_sync(0xFFFFFFFFu, tot_r, o);
            }
            if (lane == 0) {
                if (tot_f) atomicAdd(J.totals, (unsigned long long)tot_f);
                if (tot_r) atomicAdd(J.totals + 1, (unsigned long long)tot_r);
            }
            tot_f = tot_r = 0;
        }
        for (uint32_t i = threadIdx.x; i < NC * 32 * EV_HIST_STRIDE; i += blockDim.x) {
            const uint32_t v = hist[i];
            if (!v) continue;
            hist[i] = 0;
            const uint32_t n = i / (32 * EV_HIST_STRIDE);
            const uint32_t r = i % (32 * EV_HIST_STRIDE);
            const uint32_t c = r / EV_HIST_STRIDE, b = r % EV_HIST_STRIDE;
            if (b >= 32 || c >= P.n_cols) continue;
            const uint32_t k = 32 * (P.col0 + c) + b;     // window bit index
            unsigned long long* dst;
            if (KIND == EV_FWD) dst = J.counters + (size_t)(n == 0 ? C_NCC : n == 1 ? C_MFWD : C_MSCC) * P.stride;
            else if (KIND == EV_REV) dst = J.counters + (size_t)C_MREV * P.stride;
            else dst = J.aux + (size_t)n * P.stride;
            atomicAdd(dst + k, (unsigned long long)v);
        }
        __syncthreads();
    };

    // The CTA's tile range is cut into job segments; inside a segment warp w owns the tiles
    // seg_lo + w, + EV_WARPS, ... (neighbouring warps read neighbouring memory) and the warps
    // only meet at the segment end, where the counters of the job are flushed.
    uint32_t itw = 0;                         // tiles this warp has consumed (stage = itw & 1)
    uint32_t job = 0;
    while (job + 1 < P.n_jobs && P.jobs[job + 1].tile_begin <= t_lo) job++;
    for (uint32_t seg_lo = t_lo; seg_lo < t_hi; ++job) {
        const EventJob& J = P.jobs[job];
        const uint32_t seg_hi = min(t_hi, J.tile_begin + J.n_tiles);
        if (!job_active(J)) { seg_lo = seg_hi; continue; }
        // the job's fields in registers: the mbarrier / TMA asm clobbers memory, so every J.x in
        // the loops below would be a fresh global load
        const uint32_t J_words = J.array_words, J_tile_begin = J.tile_begin;
        const long long J_lo = J.a_lo, J_hi = J.a_hi;
        const int J_edge_off = J.edge_off;
        const long long N = (long long)J_words * 32;
        const bool job_ncc = NCC && (J.flags & EVF_NCC);
        const bool job_mscc = MSCC && (J.flags & EVF_MSCC);
        if (lane == 0) {
            if (seg_lo + warp < seg_hi) request(seg_lo + warp, job, itw & 1);
            if (seg_lo + warp + EV_WARPS < seg_hi) request(seg_lo + warp + EV_WARPS, job, (itw + 1) & 1);
        }
      for (uint32_t t = seg_lo + warp; t < seg_hi; t += EV_WARPS, ++itw) {
        const uint32_t g0 = (t - J_tile_begin) * EV_TILE;
        // most tiles lie wholly inside the counted range: no clip masks then (64-bit compares)
        const bool tile_inside = ((long long)g0 * 32 - 32 >= J_lo) && (((long long)g0 + EV_TILE) * 32 <= J_hi);
        // REV: every shift of every event of the tile has its j = q - k inside the counted range
        const bool tile_full = ((long long)g0 * 32 - J_lo >= D) && (((long long)g0 + EV_TILE) * 32 <= J_hi);
        auto clip = [&](uint32_t word_index) -> uint32_t {
            return tile_inside ? 0xFFFFFFFFu : range_mask((long long)word_index * 32, J_lo, J_hi);
        };
        const int s = itw & 1;
        const uint32_t* st = stage0 + (size_t)s * STAGE;
        const long long ra = region_a(J, g0);
        const long long rb = (KIND == EV_REV) ? region_b(J, g0) : 0;
        mbar_wait(&full[s], (itw >> 1) & 1);

        for (uint32_t gi = 0; gi < EV_TILE / 32; ++gi) {
            const uint32_t g = g0 + 32 * gi;
            if (g >= J_words) break;
            const uint32_t i = g + lane;
            const bool in = i < J_words;
            const long long G = (long long)g * 32;
            uint32_t ev = 0, ev2 = 0, mword = 0;     // ev: reads / run starts; ev2: run ends
            if (KIND == EV_FWD) {
                if (in) {
                    ev = st[32 * gi + lane];
                    if (job_mscc) mword = st[EV_TILE + 32 * gi + lane];
                    if (P.col0 == 0) {
                        tot_f += __popc(ev);        // F is clipped to the counted range at build
                        tot_r += __popc(st[2 * EV_TILE + 32 * gi + lane] & clip(i));
                    }
                }
            } else if (KIND == EV_REV) {
                if (in) ev = st[32 * gi + lane];
            } else if (in) {
                const uint32_t m = st[4 + 32 * gi + lane] & clip(i);
                uint32_t prev = 0;
                if (i > 0) prev = (st[3 + 32 * gi + lane] & clip(i - 1)) >> 31;
                const uint32_t up = (m << 1) | prev;     // Mc[a-1]
                ev = m & ~up;
                ev2 = ~m & up;
            }
            uint32_t nz = __ballot_sync(0xFFFFFFFFu, (ev | ev2) != 0);
            if (!nz) continue;

            // ---- this group's spans inside the staged regions
            const uint32_t *sp0, *sp1, *sp2 = nullptr;
            uint32_t r0, r1 = 0;                // bit offsets of the span origins
            long long gl = 0, gh = 0;           // REV: k-range of the shard, relative to the group
            bool full_rng = true;
            if constexpr (KIND == EV_FWD) {
                sp0 = st + 2 * EV_TILE + 32 * gi;                                   // R: start = G + e
                sp1 = st + 2 * EV_TILE + EV_REGION;                                 // M region
                // lane 0's lowest bit of M[p+rl-1-k], k = 32 col0 .. +31, for the group's e = 0
                r0 = (uint32_t)(G + rl - 32 - 32ll * P.col0 - 32 * ra);
            } else if constexpr (KIND == EV_REV) {
                if (!tile_full) {               // tiles near the ends of the counted range only
                    gl = G - J_lo;                  // kmax = min(D, gl + e)
                    gh = G - J_hi + 1;              // kmin = max(0, gh + e)
                    if (gl + 1023 < 0 || gh > D) continue;          // no event of the group counts
                    full_rng = (gh + 1023 <= 0) && (gl >= D);
                }
                sp0 = st + EV_TILE;                                                 // M region
                r0 = (uint32_t)(G - 31 - 32ll * P.col0 - 32 * ra);                  // M[q-k], as above
                const long long B4 = (N >> 1) - 1 - (G >> 1) - hmax;
                r1 = (uint32_t)(B4 & 31);
                sp1 = st + EV_TILE + EV_REGION + (uint32_t)((B4 >> 5) + P.col0 - rb);   // Er
                sp2 = sp1 + EV_REGION_HALF;                                         // Or
            } else {
                sp0 = st + EV_TILE + 4;                                             // M region
                sp1 = sp0;
                r0 = (uint32_t)(G + J_edge_off + 32ll * P.col0 - 32 * ra);          // window start, e = 0
            }

            while (nz) {
                const int src = __ffs(nz) - 1;
                nz &= nz - 1;
                uint32_t e1 = __shfl_sync(0xFFFFFFFFu, ev, src);
                uint32_t e2 = (KIND == EV_EDGE) ? __shfl_sync(0xFFFFFFFFu, ev2, src) : 0u;
                const uint32_t msrc = (KIND == EV_FWD && MSCC) ? __shfl_sync(0xFFFFFFFFu, mword, src) : 0u;
                while (e1 | e2) {
                    const bool is_end = (e1 == 0);
                    const uint32_t ecur = is_end ? e2 : e1;
                    const int bit = __ffs(ecur) - 1;
                    if (is_end) e2 = ecur & (ecur - 1);
                    else e1 = ecur & (ecur - 1);
                    uint32_t eo = 32u * src + bit;                // event offset inside the group
                    // REV saturates the uniform datapath (73 % busy, half the ALU's rate): hand the
                    // offset over in a vector register so the index arithmetic runs on the ALU
                    if constexpr (KIND == EV_REV) eo += vector_zero;
                    if constexpr (KIND == EV_FWD) {
                        sl.add_with(hist, lane, [&](uint32_t (&x)[3]) {
                            const uint32_t w1 = span_word(sp0, eo, lane) & col_mask;
                            x[0] = job_ncc ? w1 : 0u;
                            x[1] = 0u;
                            x[2] = 0u;
                            if (job_mscc && ((msrc >> bit) & 1u)) {          // M[p], warp-uniform
                                const uint32_t w2 = span_word_rev(sp1, r0 + eo, lane) & col_mask;
                                x[1] = w2;
                                x[2] = w1 & w2;
                            }
                        });
                    } else if constexpr (KIND == EV_REV) {
                        uint32_t kmask = col_mask;
                        if (!full_rng) {  // shifts k with j = q - k inside the counted range
                            long long kmin = gh + eo, kmax = gl + eo;
                            if (kmin < 0) kmin = 0;
                            if (kmax > D) kmax = D;
                            if (kmin > kmax) continue;           // warp-uniform
                            kmask &= range_mask((long long)(P.col0 + lane) * 32, kmin, kmax + 1);
                        }
                        const uint32_t tl = eo + (uint32_t)rl - 1u;              // (q + rl - 1) - G
                        const uint32_t* dec = (tl & 1u) ? sp2 : sp1;
                        sl.add_with(hist, lane, [&](uint32_t (&x)[1]) {
                            const uint32_t w3 = span_word_rev(sp0, r0 + eo, lane);
                            const uint32_t w4 = span_word(dec, r1 + (uint32_t)hmax - (tl >> 1), lane);
                            x[0] = w3 & w4 & kmask;
                        });
                    } else {
                        // run starts / ends of M inside the window: dM = M[a] - M[a-1]
                        sl.add_with(hist, lane, [&](uint32_t (&x)[2]) {
                            const uint32_t m1 = span_word(sp0, r0 + eo, lane);
                            const uint32_t m0 = span_word(sp0, r0 + eo - 1u, lane);
                            const uint32_t ws = m1 & ~m0 & col_mask;
                            const uint32_t we = ~m1 & m0 & col_mask;
                            x[0] = is_end ? we : ws;     // start.start + end.end
                            x[1] = is_end ? ws : we;     // start.end + end.start
                        });
                    }
                }
            }
        }
        __syncwarp();      // every lane is done with stage s
        if (lane == 0 && t + 2 * EV_WARPS < seg_hi) request(t + 2 * EV_WARPS, job, s);
      }
        flush_block(job);
        seg_lo = seg_hi;
    }
}

// ---------------------------------------------------------------- derive
// One streaming pass over the (unclipped) M window of a job:
//   Er/Or  = even / odd bits of M, bit-reversed (W/2 words each): the stride-2 descending window
//            M[q+rl-1-2k] of a reverse read becomes an ascending window of one of them
//   adirect[0..1] += sum_j Mc[j] M[j + lag0/lag1]   (Mc = M clipped to [a_lo, a_hi))
__device__ __forceinline__ uint32_t compress_even(uint64_t x)
{
    x &= 0x5555555555555555ull;
    x = (x | (x >> 1)) & 0x3333333333333333ull;
    x = (x | (x >> 2)) & 0x0f0f0f0f0f0f0f0full;
    x = (x | (x >> 4)) & 0x00ff00ff00ff00ffull;
    x = (x | (x >> 8)) & 0x0000ffff0000ffffull;
    x = (x | (x >> 16)) & 0x00000000ffffffffull;
    return (uint32_t)x;
}

// Even bits of x in the low half, odd bits in the high half (outer perfect unshuffle,
// Hacker's Delight 7-2; the byte-level step is one PRMT): 13 instructions for both halves.
__device__ __forceinline__ uint32_t unshuffle16(uint32_t x)
{
    uint32_t t;
    t = (x ^ (x >> 1)) & 0x22222222u; x ^= t ^ (t << 1);
    t = (x ^ (x >> 2)) & 0x0C0C0C0Cu; x ^= t ^ (t << 2);
    t = (x ^ (x >> 4)) & 0x00F000F0u; x ^= t ^ (t << 4);
    return __byte_perm(x, 0u, 0x3120);
}

// Er / Or words of one quad of M (see derive_multi_kernel): even / odd bits, bit-reversed
__device__ __forceinline__ void store_decimated(uint32_t* __restrict__ Er, uint32_t* __restrict__ Or, uint32_t at,
                                                uint32_t m0, uint32_t m1, uint32_t m2, uint32_t m3)
{
    const uint32_t u0 = unshuffle16(m0), u1 = unshuffle16(m1), u2 = unshuffle16(m2), u3 = unshuffle16(m3);
    *reinterpret_cast<uint2*>(Er + at) =
        make_uint2(__brev(__byte_perm(u2, u3, 0x5410)), __brev(__byte_perm(u0, u1, 0x5410)));
    *reinterpret_cast<uint2*>(Or + at) =
        make_uint2(__brev(__byte_perm(u2, u3, 0x7632)), __brev(__byte_perm(u0, u1, 0x7632)));
}

__device__ __forceinline__ uint32_t load_or_zero(const uint32_t* __restrict__ M, long long w, long long W)
{
    return (w >= 0 && w < W) ? __ldg(M + w) : 0u;
}

// One thread per 4 words (16-byte loads / stores; array_words is a multiple of 4); all jobs of
// the batch in ONE launch: a CTA owns DERIVE_QUADS_PER_CTA quads of one task (block -> task by
// binary search over the prefix sum of the tasks' CTA counts).
struct DeriveTask {
    const uint32_t* M;
    uint32_t* Er;
    uint32_t* Or;
    unsigned long long* adirect;
    long long a_lo, a_hi;
    long long lag0, lag1;
    uint32_t array_words;
    uint32_t n_intervals;           // mask_build_multi_kernel only, like the three below
    const uint32_t* begins;
    const uint32_t* ends;
    long long p_off;                // chromosome bit of array bit 0 (negative inside the left halo)
    unsigned long long* aux;        // edge_list_kernel: [2][stride] edge sums of the job
    int edge_off;                   //   window start relative to an edge (as EventJob::edge_off)
    int reserved;
};
constexpr uint32_t DERIVE_QUADS_PER_CTA = 2048;

__global__ void __launch_bounds__(256) derive_multi_kernel(const DeriveTask* __restrict__ tasks,
                                                           const uint64_t* __restrict__ cta_prefix,
                                                           uint32_t n_tasks, int n_lags,
                                                           const int* __restrict__ gate)
{
    if (gate && !*gate) return;      // fallback of mask_build_multi_kernel
    // `tasks` / `cta_prefix` may point into the middle of the batch's tables (one group of jobs)
    const uint64_t cta = cta_prefix[0] + blockIdx.x;
    uint32_t lo_t = 0, hi_t = n_tasks;          // cta_prefix[lo_t] <= cta < cta_prefix[hi_t]
    while (hi_t - lo_t > 1) {
        const uint32_t mid = (lo_t + hi_t) >> 1;
        if (cta_prefix[mid] <= cta) lo_t = mid; else hi_t = mid;
    }
    const DeriveTask& T = tasks[lo_t];
    const uint32_t* __restrict__ M = T.M;
    uint32_t* __restrict__ Er = T.Er;
    uint32_t* __restrict__ Or = T.Or;
    const uint32_t array_words = T.array_words;
    const long long a_lo = T.a_lo, a_hi = T.a_hi;
    const uint32_t quads = array_words >> 2;
    const uint32_t pairs = array_words >> 1;
    const long long W = array_words;
    const long long q0 = T.lag0 >> 5, q1 = T.lag1 >> 5;
    const uint32_t s0 = (uint32_t)(T.lag0 & 31), s1 = (uint32_t)(T.lag1 & 31);
    const uint32_t v_lo = (uint32_t)(cta - cta_prefix[lo_t]) * DERIVE_QUADS_PER_CTA;
    const uint32_t v_hi = min(quads, v_lo + DERIVE_QUADS_PER_CTA);
    unsigned int c0 = 0, c1 = 0;
    for (uint32_t v = v_lo + threadIdx.x; v < v_hi; v += blockDim.x) {
        const uint32_t i = 4 * v;
        const uint4 m4 = *reinterpret_cast<const uint4*>(M + i);
        const uint32_t m[4] = {m4.x, m4.y, m4.z, m4.w};
        if (Er) store_decimated(Er, Or, pairs - 2 - (i >> 1), m[0], m[1], m[2], m[3]);
        // directly counted lags: sum_j Mc[j] M[j + lag]
        uint32_t mc[4];
        uint32_t any = 0;
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            mc[t] = m[t] & range_mask((long long)(i + t) * 32, a_lo, a_hi);
            any |= mc[t];
        }
        if (any) {
            uint32_t n0[5], n1[5];
#pragma unroll
            for (int t = 0; t < 5; ++t) n0[t] = load_or_zero(M, (long long)i + q0 + t, W);
            if (n_lags > 1) {
#pragma unroll
                for (int t = 0; t < 5; ++t) n1[t] = (q1 == q0) ? n0[t] : load_or_zero(M, (long long)i + q1 + t, W);
            }
#pragma unroll
            for (int t = 0; t < 4; ++t) {
                c0 += __popc(mc[t] & __funnelshift_r(n0[t], n0[t + 1], s0));
                if (n_lags > 1) c1 += __popc(mc[t] & __funnelshift_r(n1[t], n1[t + 1], s1));
            }
        }
    }
    for (int o = 16; o; o >>= 1) {
        c0 += __shfl_xor_sync(0xFFFFFFFFu, c0, o);
        c1 += __shfl_xor_sync(0xFFFFFFFFu, c1, o);
    }
    if ((threadIdx.x & 31) == 0) {
        if (c0) atomicAdd(T.adirect, (unsigned long long)c0);
        if (c1) atomicAdd(T.adirect + 1, (unsigned long long)c1);
    }
}

// ---------------------------------------------------------------- mask build (gather)
// M, Er and Or in ONE pass, without memset, atomics or a second read of M.
// BigWig intervals are sorted and disjoint (check_intervals_multi_kernel verifies it on the
// device; otherwise the fill + derive kernels above run instead, gated on the same flag), so
// M can be GATHERED: a warp owns 32 consecutive quads (4096 bits), finds the first interval
// that reaches them (one binary search per warp, then a running cursor), loads the candidate
// intervals 32 at a time (coalesced) and every lane ORs the candidates into its own quad.
// The two directly counted lags come from the intervals themselves (check_intervals_multi_kernel).
__device__ __forceinline__ uint32_t ones_from(int lo)      // bits [lo, 32), lo in [0, inf)
{
    return __funnelshift_lc(0u, 0xFFFFFFFFu, (uint32_t)lo);
}
__device__ __forceinline__ uint32_t ones_below(int hi)     // bits [0, hi), hi in (-inf, 32]
{
    return __funnelshift_rc(0xFFFFFFFFu, 0u, (uint32_t)(32 - hi));
}

__global__ void __launch_bounds__(256) check_intervals_multi_kernel(const FillTask* __restrict__ tasks,
                                                                    const uint64_t* __restrict__ prefix,
                                                                    uint32_t n_tasks, int n_lags,
                                                                    int* __restrict__ err_flag,
                                                                    int* __restrict__ unsorted)
{
    // err_flag: an interval lies beyond the array (error, as in fill_multi_kernel)
    // unsorted: begins / ends are not sorted and disjoint -> the gather is not valid
    // Also the two directly counted lags, sum_j Mc[j] M[j + lag], straight from the intervals:
    // with disjoint sorted runs it is the overlap of run k (clipped to the counted range) with
    // run k, k+1, ... moved down by `lag` -- a handful of integer operations per interval instead
    // of a popcount pass over the whole array.  (Meaningless for unsorted input, and unused then.)
    const uint64_t total = prefix[n_tasks];
    const int lane = threadIdx.x & 31;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    // warp-uniform trip count: the lag sums are reduced over the warp before they go to memory
    for (uint64_t wb = prefix[0] + (uint64_t)blockIdx.x * blockDim.x + (threadIdx.x & ~31u); wb < total; wb += stride) {
        const uint64_t i = wb + lane;
        const bool on = i < total;
        const uint32_t t = find_task(prefix, n_tasks, on ? i : total - 1);
        const FillTask& T = tasks[t];
        unsigned long long s0 = 0, s1 = 0;
        if (on) {
            const uint64_t k = i - prefix[t], n = prefix[t + 1] - prefix[t];
            const uint32_t b = T.begins[k], e = T.ends[k];
            if (e > b && (uint64_t)e >= T.nbits) atomicExch(err_flag, 1);
            if (b > e) atomicExch(unsorted, 1);
            if (k + 1 < n && e > T.begins[k + 1]) atomicExch(unsorted, 1);
            // run k = chromosome bits b+1 .. e; its counted part A
            const long long A_lo = max(max((long long)b + 1, (long long)T.c_lo), (long long)T.p_lo);
            const long long A_hi = min(min((long long)e, (long long)T.c_hi - 1), (long long)T.p_hi - 1);
            const long long L0 = T.lag0, L1 = T.lag1, Lmax = max(L0, L1);
            long long prev = b;
            int guard = 0;
            for (uint64_t kk = k; kk < n && A_hi >= A_lo; ++kk) {
                const long long bb = T.begins[kk], ee = T.ends[kk];
                const long long B_lo = max(bb + 1, (long long)T.p_lo), B_hi = min(ee, (long long)T.p_hi - 1);
                if (bb < prev || B_lo - Lmax > A_hi) break;
                prev = bb;
                if (++guard > 4096) { atomicExch(unsorted, 1); break; }   // absurd input: let the fallback count
                if (B_hi < B_lo) continue;
                const long long o0 = min(A_hi, B_hi - L0) - max(A_lo, B_lo - L0) + 1;
                if (o0 > 0) s0 += (unsigned long long)o0;
                const long long o1 = min(A_hi, B_hi - L1) - max(A_lo, B_lo - L1) + 1;
                if (n_lags > 1 && o1 > 0) s1 += (unsigned long long)o1;
            }
        }
        const uint32_t t0 = __shfl_sync(0xFFFFFFFFu, t, 0);
        if (__all_sync(0xFFFFFFFFu, t == t0)) {         // one task in the warp: one atomic pair
            for (int o = 16; o; o >>= 1) {
                s0 += __shfl_xor_sync(0xFFFFFFFFu, s0, o);
                s1 += __shfl_xor_sync(0xFFFFFFFFu, s1, o);
            }
            if (lane == 0) {
                if (s0) atomicAdd(T.agather, s0);
                if (s1) atomicAdd(T.agather + 1, s1);
            }
        } else {
            if (s0) atomicAdd(T.agather, s0);
            if (s1) atomicAdd(T.agather + 1, s1);
        }
    }
}

__global__ void __launch_bounds__(256) mask_build_multi_kernel(const DeriveTask* __restrict__ tasks,
                                                               const uint64_t* __restrict__ cta_prefix,
                                                               uint32_t n_tasks,
                                                               const int* __restrict__ unsorted)
{
    if (*unsorted) return;                      // the scatter-style fallback builds M instead
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint64_t cta = cta_prefix[0] + blockIdx.x;
    uint32_t lo_t = 0, hi_t = n_tasks;
    while (hi_t - lo_t > 1) {
        const uint32_t mid = (lo_t + hi_t) >> 1;
        if (cta_prefix[mid] <= cta) lo_t = mid; else hi_t = mid;
    }
    const DeriveTask& T = tasks[lo_t];
    uint32_t* __restrict__ M = const_cast<uint32_t*>(T.M);
    uint32_t* __restrict__ Er = T.Er;
    uint32_t* __restrict__ Or = T.Or;
    const uint32_t* __restrict__ begins = T.begins;
    const uint32_t* __restrict__ ends = T.ends;
    const uint32_t nI = T.n_intervals;
    const long long p_off = T.p_off;
    const uint32_t quads = T.array_words >> 2;
    const uint32_t pairs = T.array_words >> 1;
    const uint32_t v_lo = (uint32_t)(cta - cta_prefix[lo_t]) * DERIVE_QUADS_PER_CTA;
    const uint32_t v_hi = min(quads, v_lo + DERIVE_QUADS_PER_CTA);
    const uint32_t per_warp = DERIVE_QUADS_PER_CTA / 8;
    const uint32_t w_lo = v_lo + warp * per_warp;
    const uint32_t w_hi = min(v_hi, w_lo + per_warp);
    const uint32_t step = 32;
    const long long POS = 1ll << 40;

    // first interval whose end reaches the warp's first bit: a 32-ary search (one coalesced
    // probe per round, 3-4 rounds) instead of 17 dependent loads of a binary search
    uint32_t idx = 0;                           // first interval the NEXT group may need
    if (w_lo < w_hi) {
        const long long P0 = (long long)w_lo * 128 + p_off;
        uint32_t lo = 0, hi = nI;               // the answer lies in [lo, hi]
        while (hi - lo >= 33) {
            const uint32_t width = hi - lo;
            const uint32_t pi = lo + (uint32_t)(((uint64_t)width * (lane + 1)) / 33);
            const uint32_t cnt = __popc(__ballot_sync(0xFFFFFFFFu, (long long)ends[pi] < P0));   // a prefix
            const uint32_t nlo = cnt ? lo + (uint32_t)(((uint64_t)width * cnt) / 33) + 1 : lo;
            const uint32_t nhi = cnt < 32 ? lo + (uint32_t)(((uint64_t)width * (cnt + 1)) / 33) : hi;
            lo = nlo; hi = nhi;
        }
        const uint32_t pi = lo + lane;
        idx = lo + __popc(__ballot_sync(0xFFFFFFFFu, pi < hi && (long long)ends[pi] < P0));
    }
    bool have = false;                          // (cb, ce) hold the candidates c_loaded .. c_loaded+31
    uint32_t c_loaded = 0;
    long long cb = POS, ce = POS;
    for (uint32_t base = w_lo; base < w_hi; base += step) {
        const uint32_t v = base + lane;
        const long long G_lo = (long long)base * 128 + p_off;      // chromosome bit of the group's bit 0
        const long long G_next = G_lo + 128ll * step;              // ... of the next group's
        uint32_t m0 = 0, m1 = 0, m2 = 0, m3 = 0;
        // ~13 intervals reach a warp's 32768 bits: the batch in registers usually serves every group
        uint32_t c = (have && idx >= c_loaded && idx < c_loaded + 32) ? c_loaded : idx;
        bool front = true;                      // every candidate before c ends before G_next
        for (;; c += 32) {
            if (!have || c != c_loaded) {
                const uint32_t ci = c + lane;
                cb = POS; ce = POS;                                 // sentinels past the last interval
                if (ci < nI) { cb = begins[ci]; ce = ends[ci]; }
                have = true;
                c_loaded = c;
            }
            // this lane's candidate as bit offsets into the group: bits [rb, re] (chromosome b+1 .. e)
            const unsigned below = __ballot_sync(0xFFFFFFFFu, ce < G_lo);             // a prefix of the lanes
            const unsigned beyond = __ballot_sync(0xFFFFFFFFu, cb + 1 >= G_lo + 4096);  // a suffix
            if (front) {
                const unsigned done = __ballot_sync(0xFFFFFFFFu, ce < G_next);         // a prefix, too
                idx = c + __popc(done);
                front = (done == 0xFFFFFFFFu);
            }
            const int rb = (int)max(cb + 1 - G_lo, 0ll), re = (int)min(ce - G_lo, 4095ll);
            unsigned live = ~below & ~beyond;
            while (live) {
                const int src = __ffs(live) - 1;
                live &= live - 1;
                const int s = __shfl_sync(0xFFFFFFFFu, rb, src) - 128 * lane;   // relative to the quad
                const int t = __shfl_sync(0xFFFFFFFFu, re, src) - 128 * lane;
                if (s <= t && t >= 0 && s <= 127) {
                    m0 |= ones_from(max(s, 0)) & ones_below(min(t + 1, 32));
                    m1 |= ones_from(max(s - 32, 0)) & ones_below(min(t + 1 - 32, 32));
                    m2 |= ones_from(max(s - 64, 0)) & ones_below(min(t + 1 - 64, 32));
                    m3 |= ones_from(max(s - 96, 0)) & ones_below(min(t + 1 - 96, 32));
                }
            }
            if (beyond) break;                  // sentinels (past the last interval) count as beyond
        }
        if (v < w_hi) {                         // (w_hi <= quads: nothing exists past the array)
            const uint32_t i = 4 * v;
            *reinterpret_cast<uint4*>(M + i) = make_uint4(m0, m1, m2, m3);
            if (Er) store_decimated(Er, Or, pairs - 2 - (i >> 1), m0, m1, m2, m3);
        }
    }
}

// ---------------------------------------------------------------- run edges from the interval list
// With sorted, disjoint intervals (the gather build's precondition) the run edges of Mc are
// KNOWN: interval (b, e] clipped to the counted range starts a run at s and ends one at t.  So
// the edge sums need no scan over M at all: a CTA takes 256 consecutive intervals of one job, a
// warp 32 of them, and every edge adds its window of dM straight from global memory (two
// coalesced words per lane, the next interval's words in flight while this one is counted).
// Touching and empty intervals produce a start and an end at the same bit, whose windows cancel
// exactly in C = plus - minus.  The scan-based run-edge kernel stays behind this one as the gated
// fallback for unsorted input.
constexpr uint32_t EDGE_INTERVALS_PER_CTA = 256;

struct EdgeListParams {
    uint32_t col0, n_cols, max_shift, stride;
};

__global__ void __launch_bounds__(256) edge_list_kernel(const DeriveTask* __restrict__ tasks,
                                                        const uint64_t* __restrict__ cta_prefix,
                                                        uint32_t n_tasks, EdgeListParams P,
                                                        const int* __restrict__ unsorted)
{
    if (*unsorted) return;
    __shared__ uint32_t hist[2 * 32 * EV_HIST_STRIDE];
    const int lane = threadIdx.x & 31;
    const int warp = __shfl_sync(0xFFFFFFFFu, (int)(threadIdx.x >> 5), 0);
    const uint64_t cta = cta_prefix[0] + blockIdx.x;
    uint32_t lo_t = 0, hi_t = n_tasks;
    while (hi_t - lo_t > 1) {
        const uint32_t mid = (lo_t + hi_t) >> 1;
        if (cta_prefix[mid] <= cta) lo_t = mid; else hi_t = mid;
    }
    const DeriveTask& T = tasks[lo_t];
    const uint32_t* __restrict__ M = T.M;
    const uint32_t nI = T.n_intervals;
    for (uint32_t i = threadIdx.x; i < 2 * 32 * EV_HIST_STRIDE; i += blockDim.x) hist[i] = 0;
    __syncthreads();

    // this lane's interval as array bits: a run of Mc starts at s and ends at t (first bit after
    // it).  What travels through the warp is the window start o = edge + edge_off + 32 col0 of
    // each edge as ONE 32-bit value (the host only takes this path for arrays under 2^32 bits; the
    // halo keeps every counted window, and the bit below it, inside the array).
    const uint32_t iv = (uint32_t)(cta - cta_prefix[lo_t]) * EDGE_INTERVALS_PER_CTA + 32u * warp + lane;
    const long long shift0 = (long long)T.edge_off + 32ll * P.col0;     // window start relative to the edge
    uint32_t os = 0, ot = 0;
    bool valid = false;
    if (iv < nI) {
        const long long b = T.begins[iv], e = T.ends[iv];
        const long long s = max(b + 1 - T.p_off, T.a_lo), t = min(e + 1 - T.p_off, T.a_hi);
        valid = s < t;
        os = (uint32_t)(s + shift0);
        ot = (uint32_t)(t + shift0);
    }
    const unsigned vmask = __ballot_sync(0xFFFFFFFFu, valid);
    const uint32_t col_mask = lane < (int)P.n_cols ? 0xFFFFFFFFu : 0u;
    const uint32_t w_last = T.array_words - 1;      // columns past max_shift may point past the array: clamp (masked)

    // the window words of lane `lane` for a window starting at array bit o
    auto fetch = [&](uint32_t o, uint32_t& w0, uint32_t& w1, uint32_t& wm) {
        const uint32_t wi = (o >> 5) + lane;
        w0 = __ldg(M + min(wi, w_last));
        w1 = __ldg(M + min(wi + 1, w_last));
        wm = 0u;
        if ((o & 31u) == 0u) wm = __ldg(M + min(wi - 1, w_last));      // bit o-1 (uniform branch)
    };
    Slices<2> sl;
    sl.reset();
    auto count = [&](uint32_t o, bool is_end, uint32_t w0, uint32_t w1, uint32_t wm) {
        const uint32_t sh = o & 31u;
        sl.add_with(hist, lane, [&](uint32_t (&x)[2]) {
            const uint32_t m1 = __funnelshift_r(w0, w1, sh);
            const uint32_t m0 = sh ? __funnelshift_r(w0, w1, sh - 1u) : __funnelshift_r(wm, w0, 31u);
            const uint32_t ws = m1 & ~m0 & col_mask;
            const uint32_t we = ~m1 & m0 & col_mask;
            x[0] = is_end ? we : ws;     // start.start + end.end
            x[1] = is_end ? ws : we;     // start.end + end.start
        });
    };
    // software pipeline over the warp's valid intervals: the words of interval k+1 are requested
    // before interval k is counted
    unsigned todo = vmask;
    uint32_t a0 = 0, a1 = 0, am = 0, b0 = 0, b1 = 0, bm = 0;        // start / end edge of the current interval
    uint32_t cs = 0, ct = 0;
    if (todo) {
        const int k = __ffs(todo) - 1;
        todo &= todo - 1;
        cs = __shfl_sync(0xFFFFFFFFu, os, k);
        ct = __shfl_sync(0xFFFFFFFFu, ot, k);
        fetch(cs, a0, a1, am);
        fetch(ct, b0, b1, bm);
        for (;;) {
            uint32_t n0 = 0, n1 = 0, nm = 0, q0 = 0, q1 = 0, qm = 0, ns = 0, nt = 0;
            const bool more = todo != 0;
            if (more) {
                const int k2 = __ffs(todo) - 1;
                todo &= todo - 1;
                ns = __shfl_sync(0xFFFFFFFFu, os, k2);
                nt = __shfl_sync(0xFFFFFFFFu, ot, k2);
                fetch(ns, n0, n1, nm);
                fetch(nt, q0, q1, qm);
            }
            count(cs, false, a0, a1, am);
            count(ct, true, b0, b1, bm);
            if (!more) break;
            cs = ns; ct = nt;
            a0 = n0; a1 = n1; am = nm; b0 = q0; b1 = q1; bm = qm;
        }
    }
    sl.flush(hist, lane);
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < 2 * 32 * EV_HIST_STRIDE; i += blockDim.x) {
        const uint32_t v = hist[i];
        if (!v) continue;
        const uint32_t n = i / (32 * EV_HIST_STRIDE);
        const uint32_t r = i % (32 * EV_HIST_STRIDE);
        const uint32_t c = r / EV_HIST_STRIDE, b = r % EV_HIST_STRIDE;
        if (b >= 32 || c >= P.n_cols) continue;
        const uint32_t k = 32 * (P.col0 + c) + b;
        atomicAdd(T.aux + (size_t)n * P.stride + k, (unsigned long long)v);
    }
}

// ---------------------------------------------------------------- finalize
// dmp from the two direct lags and the edge products C[b] = aux[0][b] - aux[1][b]:
//   MSCC      dmp[0] = A[rl-1], dmp[1] = A[rl-2], dmp[k+1] = 2 dmp[k] - dmp[k-1] - C[D-k]
//   AUTOCORR  out[0] = A[0],    out[1] = A[1],    out[s+1] = 2 out[s] - out[s-1] - C[s]
// The recurrence is a double prefix sum (first differences d[k+1] = d[k] - C[.], then
// dmp[k] = dmp[k-1] + d[k]): one CTA per job runs two in-place block scans over the output vector.
__device__ __forceinline__ void block_scan_inplace(long long* a, uint32_t n, long long* part)
{
    const uint32_t chunk = (n + blockDim.x - 1) / blockDim.x;
    const uint32_t lo = min(n, threadIdx.x * chunk), hi = min(n, lo + chunk);
    long long sum = 0;
    for (uint32_t i = lo; i < hi; ++i) sum += a[i];
    part[threadIdx.x] = sum;
    __syncthreads();
    if (threadIdx.x == 0) {
        long long run = 0;
        for (uint32_t t = 0; t < blockDim.x; ++t) { const long long v = part[t]; part[t] = run; run += v; }
    }
    __syncthreads();
    long long run = part[threadIdx.x];
    for (uint32_t i = lo; i < hi; ++i) { run += a[i]; a[i] = run; }
    __syncthreads();
}

__global__ void __launch_bounds__(256) finalize_dmp_kernel(const EventJob* __restrict__ jobs, uint32_t n_jobs,
                                                           uint32_t stride, uint32_t max_shift,
                                                           const int* __restrict__ gates, int per_group)
{
    __shared__ long long part[256];
    const EventJob& J = jobs[blockIdx.x];
    if (!(J.flags & (EVF_MSCC | EVF_AUTOCORR))) return;       // uniform per CTA
    const bool autocorr = (J.flags & EVF_AUTOCORR) != 0;
    long long* out = (long long*)(J.counters + (size_t)C_DMP * stride);
    const long long* plus = (const long long*)J.aux;
    const long long* minus = plus + stride;
    // gates: the gather build's "unsorted" flags (one per H2D group, or flag 0 for a whole-batch pass)
    const bool gathered = gates && !gates[per_group ? J.gate_index : 0];
    const long long a0 = (long long)(gathered ? J.agather[0] : J.adirect[0]);
    const long long a1 = (long long)(gathered ? J.agather[1] : J.adirect[1]);
    if (max_shift == 0) { if (threadIdx.x == 0) out[0] = a0; return; }
    // out[0] = A0, out[1] = d[1] = A1 - A0, out[k+1] = -C[b(k)] for k = 1 .. max_shift-1
    for (uint32_t k = threadIdx.x; k <= max_shift; k += blockDim.x) {
        long long v;
        if (k == 0) v = a0;
        else if (k == 1) v = a1 - a0;
        else { const uint32_t b = autocorr ? k - 1 : max_shift - (k - 1); v = -(plus[b] - minus[b]); }
        out[k] = v;
    }
    __syncthreads();
    block_scan_inplace(out + 1, max_shift, part);        // first differences d[1..max_shift]
    block_scan_inplace(out, max_shift + 1, part);        // dmp[0..max_shift]
}

}  // namespace pmb
