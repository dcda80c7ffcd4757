// pmb_window.cuh -- the shift-window kernel (sm_100a): per-shift AND+popcount counts.
//
// Replaces the reference's per-shift loop (PyMaSC/core/bitarray/mscc.pyx:288-317), which
// for every shift k makes ~10 full passes over the chromosome's bit arrays
// (bit_array_and x3, count x3, acount x2, and two whole-array 1-bit shifts).  Here every
// (32-bit word w, shift k) pair is evaluated once, from shared memory:
//
//   Rk = funnelshift_r(R[w+q], R[w+q+1], r)        k = 32q + r       (R[j+k])
//   Mk = funnelshift_l(Mp[w-q-1], Mp[w-q], r)      Mp[x] = M[x+rl-1]  (M[j+rl-1-k])
//   D  = M[w] & Mk
//   ncc += popc(F&Rk)  dmp += popc(D)  mfwd += popc(FM&Mk)  mrev += popc(Rk&D)
//   mscc += popc(FM&Rk&Mk)                            (FM = F&M; 3-input ANDs are one LOP3)
//
// Mapping.  lane <-> shift: a warp task is (unit u, part p) = 128 consecutive shifts
// k = 128u + 32i + lane (i = 0..3, four per lane) x one contiguous part of the tile.
// All lanes read the SAME shared-memory words (broadcast LDS.128, conflict-free) and
// differ only in the funnel-shift amount, so the inner loop has no shuffles, no
// divergence and ~0.3 LDS per (word, 32 shifts).  Per-lane counters live in registers
// for a whole task and are merged into a per-block shared-memory histogram (u32,
// one slot per counter x shift), which is flushed to the 64-bit global counters only
// when the block moves to another chromosome or finishes.
//
// Carry-save counters.  POPC issues at 1/4 of the LOP3 rate, so a plain popc per AND
// mask is POPC-pipe bound.  A counter of level 1 first folds word pairs with a
// carry-save adder (sum = a^b^c, carry = maj(a,b,c): 2 LOP3) and only popcounts the
// carry word (weight 2); level 2 folds four words (6 LOP3, 1 POPC, weight 4).  The
// residual `ones`/`twos` words are popcounted once per task.  The counts are exact.
// The level per counter is a template parameter, tuned on the device so that the ALU
// (LOP3/SHF) and POPC pipes finish together.
//
// Staging.  Tiles of T words of F/FM/M plus the (T + halo) windows of R and Mp are
// brought in with 1-D TMA bulk copies (cp.async.bulk.shared::cluster.global, SASS
// UBLKCP) completing on an mbarrier, double buffered: tile t+2 is requested as soon as
// tile t has been consumed.  The kernel is persistent: gridDim.x blocks each own a
// contiguous range of the batch's tile list (all chromosomes of the batch in one launch).
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>

namespace pmb {

enum { C_NCC = 0, C_DMP = 1, C_MFWD = 2, C_MREV = 3, C_MSCC = 4 };
enum { MODE_NCC = 1, MODE_MSCC = 2, MODE_BOTH = 3 };

struct WindowJob {
    const uint32_t* F;    // body word 0 of each array (origin word of the chromosome)
    const uint32_t* FM;
    const uint32_t* M;
    const uint32_t* R;
    const uint32_t* Mp;
    unsigned long long* counters;   // [5][stride]
    uint32_t tile_begin;            // index of this job's first tile in the batch tile list
    uint32_t n_tiles;
};

struct WindowParams {
    const WindowJob* jobs;
    uint32_t n_jobs;
    uint32_t total_tiles;
    uint32_t tile_words;   // T (multiple of 4*n_parts)
    uint32_t n_units;      // U = ceil((max_shift+1)/128)
    uint32_t n_parts;      // parts per tile
    uint32_t n_shifts;     // S = max_shift+1
    uint32_t stride;       // global counter stride (128 * units of the whole batch)
    uint32_t halo_words;   // 4U+4 for the units of this launch
};

// -DPMB_CHECKED: device-side bounds checks -- list indices and histogram bins (pmb_lists.cuh), TMA source
// ranges of the event kernels (pmb_events.cuh) -- because compute-sanitizer
// is closed on the B200 pool ( tools/checked_build.sh runs the GPU tests against such a build).  A
// failed check leaves 1000 + its source line in the batch's error flag, which every fetch reads:
// the call then fails with "device-side check failed at line N".  Compiled out otherwise.
#ifdef PMB_CHECKED
#define PMB_CHECK(cond, flag_ptr) do { if (!(cond)) atomicMax((flag_ptr), 1000 + __LINE__); } while (0)
#else
#define PMB_CHECK(cond, flag_ptr) ((void)0)
#endif

// ---------------------------------------------------------------- PTX helpers
__device__ __forceinline__ uint32_t smem_u32(const void* p)
{
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// 1-D TMA bulk copy global -> shared, completion on an mbarrier (bytes multiple of 16,
// both addresses 16-byte aligned).
__device__ __forceinline__ void tma_load_1d(void* dst, const void* src, uint32_t bytes, uint64_t* bar)
{
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
            smem_u32(dst)),
        "l"(src), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}
__device__ __forceinline__ void fence_proxy_async()
{
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// ---------------------------------------------------------------- carry-save counters
__device__ __forceinline__ uint32_t xor3(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t r;
    asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
__device__ __forceinline__ uint32_t maj3(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t r;
    asm("lop3.b32 %0, %1, %2, %3, 0xE8;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
__device__ __forceinline__ uint32_t and3(uint32_t a, uint32_t b, uint32_t c)
{
    uint32_t r;
    asm("lop3.b32 %0, %1, %2, %3, 0x80;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}

template <int LEVEL>
struct Counter;

template <>
struct Counter<0> {   // plain: one POPC per word
    uint32_t acc;
    __device__ __forceinline__ void reset() { acc = 0; }
    __device__ __forceinline__ void add4(uint32_t a, uint32_t b, uint32_t c, uint32_t d)
    {
        acc += __popc(a) + __popc(b);
        acc += __popc(c) + __popc(d);
    }
    __device__ __forceinline__ uint32_t total() const { return acc; }
};

template <>
struct Counter<1> {   // pairs: 2 LOP3 + 1 POPC per two words
    uint32_t ones, acc2;
    __device__ __forceinline__ void reset() { ones = 0; acc2 = 0; }
    __device__ __forceinline__ void add4(uint32_t a, uint32_t b, uint32_t c, uint32_t d)
    {
        uint32_t c0 = maj3(ones, a, b);
        ones = xor3(ones, a, b);
        uint32_t c1 = maj3(ones, c, d);
        ones = xor3(ones, c, d);
        acc2 += __popc(c0) + __popc(c1);
    }
    __device__ __forceinline__ uint32_t total() const { return 2u * acc2 + __popc(ones); }
};

template <>
struct Counter<2> {   // quads: 6 LOP3 + 1 POPC per four words
    uint32_t ones, twos, acc4;
    __device__ __forceinline__ void reset() { ones = 0; twos = 0; acc4 = 0; }
    __device__ __forceinline__ void add4(uint32_t a, uint32_t b, uint32_t c, uint32_t d)
    {
        uint32_t t0 = maj3(ones, a, b);
        ones = xor3(ones, a, b);
        uint32_t t1 = maj3(ones, c, d);
        ones = xor3(ones, c, d);
        uint32_t f = maj3(twos, t0, t1);
        twos = xor3(twos, t0, t1);
        acc4 += __popc(f);
    }
    __device__ __forceinline__ uint32_t total() const
    {
        return 4u * acc4 + 2u * __popc(twos) + __popc(ones);
    }
};

// Levels of the five counters for one kernel variant.
template <int LN, int LD, int LF, int LR, int LM>
struct Levels {
    static constexpr int ncc = LN, dmp = LD, mfwd = LF, mrev = LR, mscc = LM;
};

template <int MODE, class LV>
struct GroupCounters {   // the counters of one 32-shift group of a lane
    Counter<LV::ncc> ncc;
    Counter<LV::dmp> dmp;
    Counter<LV::mfwd> mfwd;
    Counter<LV::mrev> mrev;
    Counter<LV::mscc> mscc;
    __device__ __forceinline__ void reset()
    {
        if (MODE & MODE_NCC) ncc.reset();
        if (MODE & MODE_MSCC) { dmp.reset(); mfwd.reset(); mrev.reset(); mscc.reset(); }
    }
};

// ---------------------------------------------------------------- shared memory plan
// [ mbarriers (2 x 8 B) | histogram u32[5][128*n_units] | stage 0 | stage 1 ]
// stage: F[T] FM[T] M[T] R[T+halo] Mp[halo+T]
__host__ __device__ inline uint32_t window_stage_words(uint32_t T, uint32_t halo)
{
    return 3 * T + 2 * (T + halo);
}
__host__ __device__ inline size_t window_smem_bytes(uint32_t T, uint32_t halo, uint32_t n_units)
{
    return 16 + (size_t)5 * 128 * n_units * 4 + (size_t)2 * window_stage_words(T, halo) * 4;
}

template <int MODE, class LV>
__global__ void __launch_bounds__(512, 1) shift_window_kernel(const WindowParams P)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw);                 // [2]
    uint32_t* hist = reinterpret_cast<uint32_t*>(smem_raw + 16);            // [5][stride]
    const uint32_t HS = 128 * P.n_units;   // histogram stride: the shifts of this launch
    uint32_t* stage0 = hist + 5 * HS;
    const uint32_t T = P.tile_words, H = P.halo_words;
    const uint32_t stage_words = window_stage_words(T, H);

    const int lane = threadIdx.x & 31;
    const int warp = threadIdx.x >> 5;
    const int n_warps = blockDim.x >> 5;

    // contiguous tile range of this block
    const uint32_t t_lo = (uint32_t)(((uint64_t)P.total_tiles * blockIdx.x) / gridDim.x);
    const uint32_t t_hi = (uint32_t)(((uint64_t)P.total_tiles * (blockIdx.x + 1)) / gridDim.x);
    if (t_lo >= t_hi) return;

    for (uint32_t i = threadIdx.x; i < 5 * HS; i += blockDim.x) hist[i] = 0;
    if (threadIdx.x == 0) {
        mbar_init(&full[0], 1);
        mbar_init(&full[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    // job lookup for a tile index (jobs are few; the walk is monotone)
    uint32_t job_of_lo = 0;
    while (job_of_lo + 1 < P.n_jobs && P.jobs[job_of_lo + 1].tile_begin <= t_lo) job_of_lo++;

    // producer: thread 0 requests one tile into one stage
    auto request = [&](uint32_t t, uint32_t job, int s) {
        const WindowJob& J = P.jobs[job];
        const size_t w0 = (size_t)(t - J.tile_begin) * T;
        uint32_t* sF = stage0 + (size_t)s * stage_words;
        uint32_t* sFM = sF + T;
        uint32_t* sM = sFM + T;
        uint32_t* sR = sM + T;
        uint32_t* sMp = sR + (T + H);
        uint32_t bytes = 0;
        if (MODE & MODE_NCC) bytes += T * 4;                       // F
        bytes += (T + H) * 4;                                      // R window
        if (MODE & MODE_MSCC) bytes += 2 * T * 4 + (T + H) * 4;    // FM, M, Mp window
        fence_proxy_async();
        mbar_expect_tx(&full[s], bytes);
        if (MODE & MODE_NCC) tma_load_1d(sF, J.F + w0, T * 4, &full[s]);
        tma_load_1d(sR, J.R + w0, (T + H) * 4, &full[s]);
        if (MODE & MODE_MSCC) {
            tma_load_1d(sFM, J.FM + w0, T * 4, &full[s]);
            tma_load_1d(sM, J.M + w0, T * 4, &full[s]);
            tma_load_1d(sMp, J.Mp + w0 - H, (T + H) * 4, &full[s]);
        }
    };

    uint32_t prod_job = job_of_lo;   // producer's job cursor (thread 0 only)
    auto job_for = [&](uint32_t t, uint32_t& cursor) {
        while (cursor + 1 < P.n_jobs && P.jobs[cursor + 1].tile_begin <= t) cursor++;
        return cursor;
    };
    if (threadIdx.x == 0) {
        request(t_lo, job_for(t_lo, prod_job), 0);
        if (t_lo + 1 < t_hi) request(t_lo + 1, job_for(t_lo + 1, prod_job), 1);
    }

    const uint32_t n_tasks = P.n_units * P.n_parts;
    const uint32_t part_words = T / P.n_parts;
    uint32_t cur_job = job_of_lo;

    auto flush_hist = [&](uint32_t job) {
        __syncthreads();
        unsigned long long* g = P.jobs[job].counters;
        for (uint32_t i = threadIdx.x; i < 5 * HS; i += blockDim.x) {
            uint32_t v = hist[i];
            if (v) {
                atomicAdd(g + (size_t)(i / HS) * P.stride + (i % HS), (unsigned long long)v);
                hist[i] = 0;
            }
        }
        __syncthreads();
    };

    for (uint32_t t = t_lo; t < t_hi; ++t) {
        const uint32_t it = t - t_lo;
        const int s = it & 1;
        uint32_t job = cur_job;
        while (job + 1 < P.n_jobs && P.jobs[job + 1].tile_begin <= t) job++;
        if (job != cur_job) {       // block moves to the next chromosome
            flush_hist(cur_job);
            cur_job = job;
        }
        mbar_wait(&full[s], (it >> 1) & 1);

        const uint32_t* sF = stage0 + (size_t)s * stage_words;
        const uint32_t* sFM = sF + T;
        const uint32_t* sM = sFM + T;
        const uint32_t* sR = sM + T;
        const uint32_t* sMp = sR + (T + H) + H;   // points at the word of tile word 0

        for (uint32_t task = warp; task < n_tasks; task += n_warps) {
            const uint32_t u = task % P.n_units;
            const uint32_t part = task / P.n_units;
            const uint32_t q4 = 4 * u;
            const uint32_t w_begin = part * part_words;

            GroupCounters<MODE, LV> G[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) G[i].reset();

            uint4 rA = *reinterpret_cast<const uint4*>(sR + w_begin + q4);
            uint4 mB = make_uint4(0, 0, 0, 0);
            if (MODE & MODE_MSCC)
                mB = *reinterpret_cast<const uint4*>(sMp + (int)w_begin - (int)q4 - 4);

#pragma unroll 1
            for (uint32_t w = w_begin; w < w_begin + part_words; w += 4) {
                uint4 f4 = make_uint4(0, 0, 0, 0), fm4 = f4, m4 = f4, mA = f4;
                if (MODE & MODE_NCC) f4 = *reinterpret_cast<const uint4*>(sF + w);
                const uint4 rB = *reinterpret_cast<const uint4*>(sR + w + q4 + 4);
                if (MODE & MODE_MSCC) {
                    fm4 = *reinterpret_cast<const uint4*>(sFM + w);
                    m4 = *reinterpret_cast<const uint4*>(sM + w);
                    mA = *reinterpret_cast<const uint4*>(sMp + (int)w - (int)q4);
                }
                const uint32_t r[8] = {rA.x, rA.y, rA.z, rA.w, rB.x, rB.y, rB.z, rB.w};
                // mp[x+4] = Mp[w - q4 + x], x = -4 .. 3
                const uint32_t mp[8] = {mB.x, mB.y, mB.z, mB.w, mA.x, mA.y, mA.z, mA.w};
                const uint32_t f[4] = {f4.x, f4.y, f4.z, f4.w};
                const uint32_t fm[4] = {fm4.x, fm4.y, fm4.z, fm4.w};
                const uint32_t m[4] = {m4.x, m4.y, m4.z, m4.w};
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    uint32_t x0[4], x1[4], x2[4], x3[4], x4[4];
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const uint32_t rk = __funnelshift_r(r[j + i], r[j + i + 1], lane);
                        if (MODE & MODE_NCC) x0[j] = f[j] & rk;
                        if (MODE & MODE_MSCC) {
                            const uint32_t mk = __funnelshift_l(mp[4 + j - i - 1], mp[4 + j - i], lane);
                            x1[j] = m[j] & mk;
                            x2[j] = fm[j] & mk;
                            x3[j] = and3(rk, m[j], mk);
                            x4[j] = and3(fm[j], rk, mk);
                        }
                    }
                    if (MODE & MODE_NCC) G[i].ncc.add4(x0[0], x0[1], x0[2], x0[3]);
                    if (MODE & MODE_MSCC) {
                        G[i].dmp.add4(x1[0], x1[1], x1[2], x1[3]);
                        G[i].mfwd.add4(x2[0], x2[1], x2[2], x2[3]);
                        G[i].mrev.add4(x3[0], x3[1], x3[2], x3[3]);
                        G[i].mscc.add4(x4[0], x4[1], x4[2], x4[3]);
                    }
                }
                rA = rB;
                mB = mA;
            }
            // merge this task's per-lane counts into the block histogram
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const uint32_t k = 128 * u + 32 * i + lane;
                if (MODE & MODE_NCC) atomicAdd(hist + C_NCC * HS + k, G[i].ncc.total());
                if (MODE & MODE_MSCC) {
                    atomicAdd(hist + C_DMP * HS + k, G[i].dmp.total());
                    atomicAdd(hist + C_MFWD * HS + k, G[i].mfwd.total());
                    atomicAdd(hist + C_MREV * HS + k, G[i].mrev.total());
                    atomicAdd(hist + C_MSCC * HS + k, G[i].mscc.total());
                }
            }
        }
        __syncthreads();   // every warp is done with stage s
        if (threadIdx.x == 0 && t + 2 < t_hi) request(t + 2, job_for(t + 2, prod_job), s);
    }
    flush_hist(cur_job);
}

}  // namespace pmb
