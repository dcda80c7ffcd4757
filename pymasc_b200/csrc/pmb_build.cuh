// pmb_build.cuh -- bit-array build kernels (sm_100a).
//
// Device layout of one chromosome job (all five arrays share one geometry):
//
//     [ halo H words | body Wpad words | halo H words ]      32-bit words, LSB-first,
//                    ^ origin word `ow` of the chromosome     bit j of the reference's
//                                                             64-bit array == bit j here
// "array bit index" a = (chromosome bit) - 32*ow + 32*H.
//   F   forward 5' ends, only bits inside the counted range [lo, hi)
//   R   reverse 5' ends, whole window (halo included: R[j+k] is read up to hi+max_shift)
//   M   mappable positions, only bits inside [lo, hi)
//   Mp  M moved down by read_len-1 bits: Mp[x] = M[x + read_len-1], whole window
//       (so that D_k[j] = M[j] & Mp[j-k], a plain "shift by k" like R)
//   FM  F & M (written by prep)
// Halos are zero, which is exactly the reference's zero-fill when it shifts arrays
// (external/BitArray/bit_array.c:1953-2011).
#pragma once
#include <stdint.h>

namespace pmb {

struct Geometry {
    int64_t origin_bit;   // 32*ow: chromosome bit of body word 0
    int64_t halo_bits;    // 32*H
    int64_t total_bits;   // 32*(H + Wpad + H)
};

// chromosome bit -> array bit, or -1 when outside the allocated window
__device__ __forceinline__ int64_t to_array_bit(const Geometry& g, int64_t chrom_bit)
{
    int64_t a = chrom_bit - g.origin_bit + g.halo_bits;
    return (a < 0 || a >= g.total_bits) ? -1 : a;
}

// K1: one thread per read position.  bit_array_set_bit (bit_array.c:668) as an
// idempotent atomicOr: duplicates collapse exactly like mscc.pyx:388-393, :416-418.
// Positions are sorted, so neighbouring threads hit the same or adjacent words.
// Only positions with clip_lo <= p < clip_hi are set (F: the counted range;
// R: everything that falls in the window).  p >= nbits raises the error flag
// (the reference would abort on its assert).
__global__ void scatter_bits_kernel(uint32_t* __restrict__ arr, Geometry g,
                                    const uint32_t* __restrict__ pos, uint64_t n,
                                    int64_t clip_lo, int64_t clip_hi, uint64_t nbits,
                                    int* __restrict__ err_flag)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (; i < n; i += stride) {
        int64_t p = pos[i];
        if ((uint64_t)p >= nbits) { atomicExch(err_flag, 1); continue; }
        if (p < clip_lo || p >= clip_hi) continue;
        int64_t a = to_array_bit(g, p);
        if (a < 0) continue;
        atomicOr(arr + (a >> 5), 1u << (a & 31));
    }
}

// K2: one warp per mappable interval (begin, end): chromosome bits begin+1 .. end
// (mscc.pyx:343-344 -> bit_array_set_region, bit_array.c:423-479), moved down by
// `shift_down` bits (0 for M, read_len-1 for Mp) and clipped to [clip_lo, clip_hi).
// Edge words use atomicOr with the partial mask, interior words a plain store.
// Intervals longer than LONG_WORDS only write their two edge words here and are
// queued for fill_long_kernel, which spreads the interior over the whole grid.
constexpr int LONG_WORDS = 4096;

__device__ __forceinline__ bool interval_span(const Geometry& g, uint32_t begin, uint32_t end,
                                              int64_t shift_down, int64_t clip_lo, int64_t clip_hi,
                                              int64_t& a0, int64_t& a1)
{
    int64_t s = (int64_t)begin + 1 - shift_down;   // first bit (inclusive)
    int64_t e = (int64_t)end - shift_down;         // last bit (inclusive)
    if (s < clip_lo) s = clip_lo;
    if (e >= clip_hi) e = clip_hi - 1;
    if (e < s) return false;
    a0 = s - g.origin_bit + g.halo_bits;
    a1 = e - g.origin_bit + g.halo_bits;
    if (a0 < 0) a0 = 0;
    if (a1 >= g.total_bits) a1 = g.total_bits - 1;
    return a1 >= a0;
}

__global__ void fill_intervals_kernel(uint32_t* __restrict__ arr, Geometry g,
                                      const uint32_t* __restrict__ begins,
                                      const uint32_t* __restrict__ ends, uint64_t n,
                                      int64_t shift_down, int64_t clip_lo, int64_t clip_hi,
                                      uint64_t nbits, int* __restrict__ err_flag,
                                      uint32_t* __restrict__ long_list, uint32_t* __restrict__ long_count)
{
    const int lane = threadIdx.x & 31;
    uint64_t warp = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    uint64_t nwarps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
    for (uint64_t i = warp; i < n; i += nwarps) {
        uint32_t b = begins[i], e = ends[i];
        if (e <= b) continue;                       // zero-length region
        if ((uint64_t)e >= nbits) { if (lane == 0) atomicExch(err_flag, 1); continue; }
        int64_t a0, a1;
        if (!interval_span(g, b, e, shift_down, clip_lo, clip_hi, a0, a1)) continue;
        int64_t fw = a0 >> 5, lw = a1 >> 5;
        uint32_t fmask = 0xFFFFFFFFu << (a0 & 31);
        uint32_t lmask = 0xFFFFFFFFu >> (31 - (a1 & 31));
        if (fw == lw) {
            if (lane == 0) atomicOr(arr + fw, fmask & lmask);
            continue;
        }
        if (lane == 0) atomicOr(arr + fw, fmask);
        if (lane == 1) atomicOr(arr + lw, lmask);
        if (lw - fw - 1 > LONG_WORDS) {
            if (lane == 0) long_list[atomicAdd(long_count, 1u)] = (uint32_t)i;
            continue;
        }
        for (int64_t w = fw + 1 + lane; w < lw; w += 32) arr[w] = 0xFFFFFFFFu;
    }
}

__global__ void fill_long_kernel(uint32_t* __restrict__ arr, Geometry g,
                                 const uint32_t* __restrict__ begins,
                                 const uint32_t* __restrict__ ends,
                                 int64_t shift_down, int64_t clip_lo, int64_t clip_hi,
                                 const uint32_t* __restrict__ long_list,
                                 const uint32_t* __restrict__ long_count)
{
    uint32_t n = *long_count;
    uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint64_t nth = (uint64_t)gridDim.x * blockDim.x;
    for (uint32_t q = 0; q < n; q++) {
        uint32_t i = long_list[q];
        int64_t a0, a1;
        if (!interval_span(g, begins[i], ends[i], shift_down, clip_lo, clip_hi, a0, a1)) continue;
        int64_t fw = a0 >> 5, lw = a1 >> 5;
        for (int64_t w = fw + 1 + (int64_t)tid; w < lw; w += (int64_t)nth) arr[w] = 0xFFFFFFFFu;
    }
}

// prep: FM = F & M over the body, strand totals over the counted range
// (bit_array_num_bits_set, bit_array.c:962-978 -> __popc).  F is already clipped to
// the range; R is masked here so that halo bits are not counted twice across shards.
__global__ void prep_kernel(const uint32_t* __restrict__ F, const uint32_t* __restrict__ M,
                            uint32_t* __restrict__ FM, const uint32_t* __restrict__ R,
                            uint64_t body_words, int64_t origin_bit, int64_t lo, int64_t hi,
                            unsigned long long* __restrict__ totals /* [2]: fwd, rev */)
{
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    unsigned int cf = 0, cr = 0;
    for (; i < body_words; i += stride) {
        uint32_t f = F[i];
        if (M) FM[i] = f & M[i];
        cf += __popc(f);
        int64_t b0 = origin_bit + (int64_t)(i << 5);   // chromosome bit of this word's bit 0
        uint32_t mask = 0xFFFFFFFFu;
        if (b0 + 32 <= lo || b0 >= hi) mask = 0;
        else {
            if (b0 < lo) mask &= 0xFFFFFFFFu << (lo - b0);
            if (b0 + 32 > hi) mask &= 0xFFFFFFFFu >> (b0 + 32 - hi);
        }
        cr += __popc(R[i] & mask);
    }
    for (int o = 16; o; o >>= 1) {
        cf += __shfl_xor_sync(0xFFFFFFFFu, cf, o);
        cr += __shfl_xor_sync(0xFFFFFFFFu, cr, o);
    }
    if ((threadIdx.x & 31) == 0) {
        if (cf) atomicAdd(totals, (unsigned long long)cf);
        if (cr) atomicAdd(totals + 1, (unsigned long long)cr);
    }
}

}  // namespace pmb

// ---------------------------------------------------------------- batched build (all jobs, one launch)
// With 24+ chromosomes per batch the per-job launches above are launch-bound (a job's scatter
// is a few microseconds of work).  The *_multi kernels take a task table and a prefix-sum of
// the task sizes; a thread (warp) finds its task by binary search and then does exactly what
// the per-job kernel does.
namespace pmb {

struct ScatterTask {
    uint32_t* arr;
    Geometry g;
    const uint32_t* pos;
    int64_t clip_lo, clip_hi;
    uint64_t nbits;
};

struct FillTask {
    uint32_t* arr;
    Geometry g;
    const uint32_t* begins;
    const uint32_t* ends;
    int64_t shift_down, clip_lo, clip_hi;
    uint64_t nbits;
    // check_intervals_multi_kernel only: the direct lags from the intervals themselves
    int64_t c_lo, c_hi;             // counted range [c_lo, c_hi), chromosome bits
    int64_t p_lo, p_hi;             // bits the array holds, [p_lo, p_hi)
    int64_t lag0, lag1;
    unsigned long long* agather;    // [2]
};

// index of the task whose [prefix[t], prefix[t+1]) holds item i (prefix has n_tasks + 1 entries)
__device__ __forceinline__ uint32_t find_task(const uint64_t* __restrict__ prefix, uint32_t n_tasks, uint64_t i)
{
    uint32_t lo = 0, hi = n_tasks;           // invariant: prefix[lo] <= i < prefix[hi]
    while (hi - lo > 1) {
        const uint32_t mid = (lo + hi) >> 1;
        if (prefix[mid] <= i) lo = mid; else hi = mid;
    }
    return lo;
}

__global__ void scatter_multi_kernel(const ScatterTask* __restrict__ tasks, const uint64_t* __restrict__ prefix,
                                     uint32_t n_tasks, int* __restrict__ err_flag, const int* __restrict__ gate)
{
    if (gate && !*gate) return;      // fallback of pack_reads_multi_kernel
    // `tasks` / `prefix` may point into the middle of the batch's tables (one group of jobs):
    // items are numbered from prefix[0]
    const uint64_t total = prefix[n_tasks];
    uint64_t i = prefix[0] + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (; i < total; i += stride) {
        const uint32_t t = find_task(prefix, n_tasks, i);
        const ScatterTask& T = tasks[t];
        const int64_t p = T.pos[i - prefix[t]];
        if ((uint64_t)p >= T.nbits) { atomicExch(err_flag, 1); continue; }
        if (p < T.clip_lo || p >= T.clip_hi) continue;
        const int64_t a = to_array_bit(T.g, p);
        if (a < 0) continue;
        atomicOr(T.arr + (a >> 5), 1u << (a & 31));
    }
}

// long_list entries: (task << 32) | interval index
__global__ void fill_multi_kernel(const FillTask* __restrict__ tasks, const uint64_t* __restrict__ prefix,
                                  uint32_t n_tasks, int* __restrict__ err_flag,
                                  unsigned long long* __restrict__ long_list, uint32_t* __restrict__ long_count,
                                  const int* __restrict__ gate)
{
    if (gate && !*gate) return;      // fallback of the gather build: runs only when the flag is set
    const int lane = threadIdx.x & 31;
    const uint64_t total = prefix[n_tasks];
    uint64_t warp = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const uint64_t nwarps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
    for (uint64_t i = prefix[0] + warp; i < total; i += nwarps) {
        const uint32_t t = find_task(prefix, n_tasks, i);
        const FillTask& T = tasks[t];
        const uint64_t k = i - prefix[t];
        const uint32_t b = T.begins[k], e = T.ends[k];
        if (e <= b) continue;                       // zero-length region
        if ((uint64_t)e >= T.nbits) { if (lane == 0) atomicExch(err_flag, 1); continue; }
        int64_t a0, a1;
        if (!interval_span(T.g, b, e, T.shift_down, T.clip_lo, T.clip_hi, a0, a1)) continue;
        const int64_t fw = a0 >> 5, lw = a1 >> 5;
        const uint32_t fmask = 0xFFFFFFFFu << (a0 & 31);
        const uint32_t lmask = 0xFFFFFFFFu >> (31 - (a1 & 31));
        if (fw == lw) {
            if (lane == 0) atomicOr(T.arr + fw, fmask & lmask);
            continue;
        }
        if (lane == 0) atomicOr(T.arr + fw, fmask);
        if (lane == 1) atomicOr(T.arr + lw, lmask);
        if (lw - fw - 1 > LONG_WORDS) {
            if (lane == 0) long_list[atomicAdd(long_count, 1u)] = ((unsigned long long)t << 32) | (uint32_t)k;
            continue;
        }
        for (int64_t w = fw + 1 + lane; w < lw; w += 32) T.arr[w] = 0xFFFFFFFFu;
    }
}

__global__ void fill_long_multi_kernel(const FillTask* __restrict__ tasks,
                                       const unsigned long long* __restrict__ long_list,
                                       const uint32_t* __restrict__ long_count,
                                       const int* __restrict__ gate)
{
    if (gate && !*gate) return;
    const uint32_t n = *long_count;
    const uint64_t tid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t nth = (uint64_t)gridDim.x * blockDim.x;
    for (uint32_t q = 0; q < n; q++) {
        const unsigned long long ent = long_list[q];
        const FillTask& T = tasks[(uint32_t)(ent >> 32)];
        const uint32_t k = (uint32_t)ent;
        int64_t a0, a1;
        if (!interval_span(T.g, T.begins[k], T.ends[k], T.shift_down, T.clip_lo, T.clip_hi, a0, a1)) continue;
        const int64_t fw = a0 >> 5, lw = a1 >> 5;
        for (int64_t w = fw + 1 + (int64_t)tid; w < lw; w += (int64_t)nth) T.arr[w] = 0xFFFFFFFFu;
    }
}

// Zero up to four buffers in one launch (16-byte stores): four cudaMemsetAsync calls cost
// ~120 us of mostly launch gaps per execute, one grid-stride pass over the same bytes ~35 us.
struct ZeroSegs {
    uint4* p[4];
    size_t n16[4];      // 16-byte units
};
__global__ void zero_segments_kernel(ZeroSegs z)
{
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
#pragma unroll
    for (int s = 0; s < 4; ++s)
        for (size_t i = tid; i < z.n16[s]; i += stride) z.p[s][i] = make_uint4(0u, 0u, 0u, 0u);
}

// zero a word range (16-byte stores), only when *gate is set
__global__ void zero_gated_kernel(uint4* __restrict__ p, size_t n16, const int* __restrict__ gate)
{
    if (gate && !*gate) return;
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (; i < n16; i += stride) p[i] = make_uint4(0u, 0u, 0u, 0u);
}

}  // namespace pmb

// ---------------------------------------------------------------- read arrays without memset or global atomics
// F and R from SORTED read positions (the host feeds them sorted, mscc.pyx:351-418 raises on
// anything else; the kernel verifies it and the atomicOr scatter above stays behind it as the
// gated fallback).  A warp takes a chunk of PACK_CHUNK consecutive reads and OWNS the words from
// the word of its first read up to the word of the next chunk's first read: it builds them
// PACK_PIECE words at a time in its own shared memory (zero, OR the reads in, copy out
// coalesced), so every word of the array is written exactly once -- zeros included -- and
// nothing has to be cleared beforehand.  Reads of the previous chunk that fall into the warp's
// first word belong to it too (it walks back over them); its own last reads that fall into
// the next chunk's first word do not.
namespace pmb {

constexpr uint32_t PACK_CHUNK = 256;      // reads per warp
constexpr uint32_t PACK_PIECE = 1024;     // words of shared memory per warp
constexpr uint32_t PACK_WARPS = 8;

struct PackTask {
    uint32_t* arr;
    Geometry g;
    const uint32_t* pos;
    uint64_t n;
    int64_t clip_lo, clip_hi;
    uint64_t nbits;
    uint32_t array_words;
    uint32_t reserved;
};

// array bit of a position clamped into [0, total_bits]: monotone in p, word = bit >> 5
__device__ __forceinline__ int64_t clamped_array_bit(const Geometry& g, int64_t p)
{
    int64_t a = p - g.origin_bit + g.halo_bits;
    return a < 0 ? 0 : (a > g.total_bits ? g.total_bits : a);
}

// A run of >= PACK_LONG_GAP read-less words is not written by its owner warp (one warp zeroing
// a centromere, or a chromosome without reads, would be the tail of the whole launch): it is
// queued, and zero_runs_kernel clears all queued runs with the whole grid.
constexpr uint32_t PACK_LONG_GAP = 16 * PACK_PIECE;
struct ZeroRun {
    uint32_t task, start, count, pad;
};

__global__ void __launch_bounds__(32 * PACK_WARPS) pack_reads_multi_kernel(const PackTask* __restrict__ tasks,
                                                                           const uint64_t* __restrict__ chunk_prefix,
                                                                           uint32_t n_tasks, int* __restrict__ err_flag,
                                                                           int* __restrict__ unsorted,
                                                                           ZeroRun* __restrict__ runs,
                                                                           uint32_t* __restrict__ run_count,
                                                                           uint32_t run_cap)
{
    __shared__ __align__(16) uint32_t piece_all[PACK_WARPS][PACK_PIECE];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t* piece = piece_all[warp];
    const uint64_t chunk_id = chunk_prefix[0] + (uint64_t)blockIdx.x * PACK_WARPS + warp;
    if (chunk_id >= chunk_prefix[n_tasks]) return;            // (no block-level barrier below)
    const uint32_t t = find_task(chunk_prefix, n_tasks, chunk_id);
    const PackTask& T = tasks[t];
    const Geometry g = T.g;
    const uint32_t* __restrict__ pos = T.pos;
    const uint64_t n = T.n;
    const uint64_t c = chunk_id - chunk_prefix[t];
    const uint64_t r_lo = c * PACK_CHUNK, r_hi = min(n, r_lo + PACK_CHUNK);
    const uint32_t W = T.array_words;
    constexpr int PER_LANE = PACK_CHUNK / 32;
    const uint32_t NONE = 0xFFFFFFFFu;
    // word index and bit mask of a read (mask 0: the read sets nothing -- clipped or outside the window)
    auto place = [&](uint32_t p, uint32_t& w, uint32_t& bit) {
        const int64_t a = (int64_t)p - g.origin_bit + g.halo_bits;
        w = (uint32_t)(clamped_array_bit(g, p) >> 5);
        const bool sets = (uint64_t)p < T.nbits && (int64_t)p >= T.clip_lo && (int64_t)p < T.clip_hi && a >= 0 &&
                          a < g.total_bits;
        bit = sets ? 1u << (a & 31) : 0u;
    };
    // all loads up front (independent: one memory latency for the whole chunk)
    uint32_t rw[PER_LANE], rb[PER_LANE];
    bool bad_order = false, bad_range = false;
#pragma unroll
    for (int k = 0; k < PER_LANE; ++k) {
        const uint64_t idx = r_lo + 32u * k + lane;
        rw[k] = NONE; rb[k] = 0u;
        if (idx < r_hi) {
            const uint32_t p = pos[idx];
            bad_range |= (uint64_t)p >= T.nbits;
            bad_order |= idx + 1 < n && pos[idx + 1] < p;
            place(p, rw[k], rb[k]);
        }
    }
    uint32_t next_w = W, next_bit;                              // first read of the next chunk
    if (r_hi < n) { place(pos[r_hi], next_w, next_bit); next_w = min(next_w, W); }
    uint32_t back_w = 0, back_bit = 0;                          // the 32 reads before this chunk
    const bool has_back = c > 0 && r_lo >= (uint64_t)lane + 1;
    if (has_back) place(pos[r_lo - 1 - lane], back_w, back_bit);
    if (bad_range) atomicExch(err_flag, 1);
    if (bad_order) atomicExch(unsorted, 1);

    const uint32_t W_lo = c == 0 ? 0u : __shfl_sync(0xFFFFFFFFu, rw[0], 0);
    const uint32_t W_hi = next_w;
    if (W_hi < W_lo || (c > 0 && W_lo == NONE)) { if (lane == 0) atomicExch(unsorted, 1); return; }
    // reads of earlier chunks that fall into this warp's first word belong to it: walk back
    uint32_t first_extra = 0;
    if (c > 0) {
        uint64_t cur = r_lo;
        for (uint32_t walked = 0;; walked += 32) {
            if (walked) {       // more than 32 earlier reads in that word: keep loading (rare)
                back_w = 0; back_bit = 0;
                if (cur >= (uint64_t)lane + 1) place(pos[cur - 1 - lane], back_w, back_bit);
            }
            const bool ok = cur >= (uint64_t)lane + 1 && back_w >= W_lo;
            const unsigned m = __ballot_sync(0xFFFFFFFFu, ok);
            const uint32_t cnt = m == 0xFFFFFFFFu ? 32u : (uint32_t)__ffs(~m) - 1u;
            first_extra |= __reduce_or_sync(0xFFFFFFFFu, (uint32_t)lane < cnt && back_w == W_lo ? back_bit : 0u);
            cur -= cnt;
            if (cnt < 32) break;
            if (walked > (1u << 16)) { if (lane == 0) atomicExch(unsorted, 1); return; }   // absurd pile-up: fallback
        }
    }
    for (uint32_t wp = W_lo; wp < W_hi;) {
        // a long read-less run ahead?  queue it for zero_runs_kernel
        uint32_t mine = W_hi;
#pragma unroll
        for (int k = PER_LANE - 1; k >= 0; --k)
            if (rw[k] >= wp && rw[k] < mine) mine = rw[k];
        const uint32_t ahead = __reduce_min_sync(0xFFFFFFFFu, mine);       // next word with a read (or W_hi)
        if (ahead > wp && ahead - wp >= PACK_LONG_GAP) {
            if (lane == 0) {
                const uint32_t q = atomicAdd(run_count, 1u);
                if (q < run_cap) runs[q] = ZeroRun{t, wp, ahead - wp, 0u};
                else atomicExch(unsorted, 1);       // cannot happen with sorted reads: let the fallback build it
            }
            wp = ahead;
            continue;
        }
        // the piece starts at the 16-byte boundary below wp, so shared and global quads line up
        const uint32_t off = wp & 3u;
        const uint32_t pw = min(PACK_PIECE - off, W_hi - wp);
        const uint32_t span = off + pw, nq = (span + 3) >> 2;
        uint4* piece4 = reinterpret_cast<uint4*>(piece);
        for (uint32_t q = lane; q < nq; q += 32) piece4[q] = make_uint4(0u, 0u, 0u, 0u);
        __syncwarp();
        if (wp == W_lo && lane == 0 && first_extra) atomicOr(piece + off, first_extra);
#pragma unroll
        for (int k = 0; k < PER_LANE; ++k)
            if (rb[k] && rw[k] >= wp && rw[k] < wp + pw) atomicOr(piece + off + (rw[k] - wp), rb[k]);
        __syncwarp();
        uint32_t* base = T.arr + (wp - off);
        for (uint32_t q = lane; q < nq; q += 32) {
            const uint32_t lo = max(off, 4 * q), hi = min(span, 4 * q + 4);     // words of this quad the warp owns
            if (hi - lo == 4) *reinterpret_cast<uint4*>(base + 4 * q) = piece4[q];
            else for (uint32_t i = lo; i < hi; ++i) base[i] = piece[i];
        }
        __syncwarp();
        wp += pw;
    }
}

__global__ void zero_runs_kernel(const PackTask* __restrict__ tasks, const ZeroRun* __restrict__ runs,
                                 const uint32_t* __restrict__ run_count, uint32_t run_cap)
{
    const uint32_t n = min(*run_count, run_cap);
    const size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    const size_t nth = (size_t)gridDim.x * blockDim.x;
    for (uint32_t q = 0; q < n; q++) {
        const ZeroRun r = runs[q];
        uint32_t* dst = tasks[r.task].arr + r.start;
        for (size_t w = tid; w < r.count; w += nth) dst[w] = 0u;
    }
}

}  // namespace pmb
