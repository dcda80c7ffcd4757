// pmb_api.cu -- C ABI of libpymasc_b200.so (see include/pymasc_b200.h).
// Host-side planning + launches for the bit-array build and shift-window kernels.
// sm_100a only; there is no CPU path in this library.
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nccl.h>      // types only: the library is bound at run time (nccl_api below), not linked
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <vector>

#include "../../include/pymasc_b200.h"
#include "pmb_build.cuh"
#include "pmb_probe.cuh"
#include "pmb_window.cuh"
#include "pmb_events.cuh"
#include "pmb_lists.cuh"

using namespace pmb;

static thread_local std::string g_create_error;

struct pmb_ctx {
    int device = 0;
    int num_sms = 148;
    cudaStream_t stream = nullptr;
    cudaStream_t own_stream = nullptr;   // created by pmb_create; `stream` may be borrowed
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    cudaEvent_t evk[2] = {nullptr, nullptr};   // event path: after the fwd / rev kernels
    std::string err;
    pmb_timing timing{};
    // options
    int variant = -1;        // -1: default per mode
    int tile_words = 0;      // 0: automatic
    int grid_blocks = 0;     // 0: one per SM
    int block_warps = 16;
    int path = 0;            // 0 auto, 1 dense shift-window, 2 event-window, 3 lists
    int list_blocks_per_sm = 0;
    int list_concurrent = -1;    // reverse + edge list kernels beside the forward kernel on the side stream: -1 auto, 0, 1
    bool timing_split = true;    // the last execute recorded the events between its kernels
    int kernel_timing = 0;       // list path: 1 = events between the kernels of an execute (pmb_get_timing's per-kernel split; they
                                 // keep a kernel from starting under the tail of the one before: +8 us per step); 0: start and end only
    int list_pair_cap = 0;       // list path: forward reads between flushes of the packed pair histogram (0: 65535; tests shrink it)
    int list_edge_items = 0;     // list path: intervals per CTA of the run-edge kernel at least (0: 512)
    int check_selftest = 0;      // -DPMB_CHECKED builds: make every TMA range check of the event kernels fail (proves the chain)
    bool event_attr_done = false; // the event kernels' shared-memory limit is raised on THIS device
    bool list_attr_done = false; // the list kernels' shared-memory limit is raised on THIS device
    ncclComm_t comm = nullptr;   // pmb_comm_init / pmb_comm_init_all
    int comm_rank = 0, comm_world = 1;
    unsigned long long* noalloc_scratch = nullptr;   // device address of g_noalloc_scratch on this device
    int list_pairs_limit = 0;    // auto: lists while the expected reverse reads per forward read and 1024 shifts stay below this
    int event_tile_words = 0;
    int event_blocks_per_sm = 0;
    int async_h2d = 0;       // prepare returns before the input copy has finished (see header)
    int64_t h2d_min_group_bytes = 16 << 20;   // no H2D group smaller than this on average
    int h2d_growth_x10 = 13; // size ratio of consecutive H2D groups, in tenths
    int h2d_groups = 4;      // event path: job groups whose build + kernels overlap the H2D of the next
    cudaStream_t copy_stream = nullptr;
    cudaStream_t side_stream = nullptr;   // grouped execute: odd groups run here, so that one group's
                                          // kernel tails overlap the next group's build
    cudaEvent_t ev_alloc = nullptr, ev_side = nullptr;
    // pinned staging for counters
    void* h_stage = nullptr;
    size_t h_stage_bytes = 0;
};

#define CU_TRY(ctx, call)                                                                   \
    do {                                                                                    \
        cudaError_t e__ = (call);                                                           \
        if (e__ != cudaSuccess) {                                                           \
            char buf__[512];                                                                \
            snprintf(buf__, sizeof buf__, "%s failed: %s (%s:%d)", #call,                   \
                     cudaGetErrorString(e__), __FILE__, __LINE__);                          \
            (ctx)->err = buf__;                                                             \
            return (e__ == cudaErrorMemoryAllocation) ? PMB_ERR_NOMEM : PMB_ERR_CUDA;       \
        }                                                                                   \
    } while (0)

// Nothing may propagate out of the C ABI: host allocation failures (std::vector, new) become
// PMB_ERR_NOMEM, anything else unexpected PMB_ERR_ARG with the text in pmb_last_error.
template <class F>
static int guarded(pmb_ctx* c, F&& body)
{
    try {
        return body();
    } catch (const std::bad_alloc&) {
        if (c) c->err = "host memory allocation failed";
        return PMB_ERR_NOMEM;
    } catch (const std::exception& e) {
        if (c) c->err = std::string("unexpected error: ") + e.what();
        return PMB_ERR_ARG;
    }
}

struct JobPlan {
    pmb_chrom_args a;          // host pointers are only valid during prepare
    uint64_t nbits = 0;
    int64_t lo = 0, hi = 0;
    int64_t origin_word = 0;   // ow
    uint64_t body_words = 0;   // Wpad = n_tiles * T
    uint32_t n_tiles = 0;
    int mode = 0;              // MODE_NCC / MODE_MSCC / MODE_BOTH; autocorr -> MODE_NCC on M
    bool autocorr = false;
    int owner = -1;            // PMB_FLAG_SAME_TRACK: the earlier job of the batch whose M this one reads
    // offsets (in uint32 words) into the batch's device buffers
    size_t in_fwd = 0, in_rev = 0, in_mb = 0, in_me = 0;
    size_t bits = 0;           // start of this job's arrays; each array_words long
    size_t bits_m = 0;         // event path: start of M (after all F/R)
    size_t bits_d = 0;         // event path: start of the derived Er|Or array (after all M)
    size_t array_words = 0;    // 2H + Wpad
};

struct WinLaunch {
    int mode;
    WindowParams P;
    uint32_t blocks, warps;
    size_t smem;
    uint64_t pairs;   // 32-bit word x shift pairs evaluated by this launch
};

struct pmb_batch {
    std::vector<JobPlan> jobs;
    std::vector<WinLaunch> launches;
    uint32_t read_len = 0, max_shift = 0, n_shifts = 0;
    uint32_t n_units = 0, stride = 0, halo_words = 0, tile_words = 0;
    uint32_t* d_in = nullptr;  size_t in_words = 0;
    uint32_t* d_bits = nullptr; size_t bits_words = 0;
    uint32_t* d_bits_raw = nullptr;   // allocation; d_bits = raw + EV_PAD_WORDS (TMA regions may overhang)
    unsigned long long* d_counters = nullptr; size_t counter_words = 0;  // + totals at the end
    WindowJob* d_jobs = nullptr; size_t d_jobs_count = 0;
    int* d_err = nullptr;
    uint32_t* d_long = nullptr; size_t long_cap = 0;   // [0] = count, [1..] = list
    // event-window path
    int path = 1;                       // 1 dense, 2 events, 3 lists
    // list path (pmb_lists.cuh): job table, entry prefixes (fwd | rev | intervals, n + 1 each),
    // difference histograms [n][4][stride], edge products [n][2][stride] + direct lags [n][2]
    bool list_ok = false;               // the inputs qualify for the list path
    size_t list_acc_copy = 0;           // words of ONE copy of the list accumulators (there are two, see ListJob)
    uint32_t list_execs = 0;            // executes so far on the list path: execute k uses copy k & 1
    bool reduce_checked = false;        // the first all-reduce of the batch has agreed on the ranks' status words
    bool list_checked = false;          // an execute on the list path has found the inputs sorted
    ListJob* d_listjobs = nullptr;
    unsigned long long* d_list_prefix = nullptr;
    unsigned long long* d_list_acc = nullptr; size_t list_acc_words = 0;
    std::vector<unsigned long long> h_list_prefix;
    std::vector<ListJob> h_listjobs;
    uint64_t total_words = 0;           // body words of all jobs (tile-size choice of the dense path)
    bool arrays_planned = false;        // the bit arrays of the dense / event path exist (layout + plan done)
    unsigned long long* d_aux = nullptr; size_t aux_words = 0;   // [n][2][stride] + [n][2]
    EventJob* d_evjobs = nullptr;
    uint32_t ev_total_tiles = 0, ev_tile_words = 0;
    bool any_ncc = false, any_mscc = false, any_autocorr = false;
    size_t zero_words = 0;              // event path: [0, zero_words) of d_bits is zeroed per execute
    size_t m_begin = 0, m_end = 0;      // event path: the M arrays of all jobs, [m_begin, m_end) of d_bits
    bool gather_m = false;              // M is gathered (mask_build_multi_kernel), not scattered
    bool edge_list_ok = false;          // every array is shorter than 2^32 bits (edge_list_kernel's 32-bit offsets)
    bool pack_reads = false;            // F, R are packed from the sorted reads (pack_reads_multi_kernel)
    uint64_t* d_edge_prefix = nullptr; uint64_t edge_ctas = 0;      // edge_list_kernel, over the derive tasks
    PackTask* d_pack = nullptr; uint64_t* d_pack_prefix = nullptr; uint32_t n_pack = 0; uint64_t pack_chunks = 0;
    ZeroRun* d_runs = nullptr; size_t runs_cap = 0;     // main | side queue
    uint32_t gather_k = 1;              // look-ahead quads of the gather
    void* d_tasks = nullptr;            // batched build: task tables + prefix sums (one allocation)
    ScatterTask* d_scatter = nullptr; uint64_t* d_scatter_prefix = nullptr; uint32_t n_scatter = 0; uint64_t scatter_items = 0;
    FillTask* d_fill = nullptr; uint64_t* d_fill_prefix = nullptr; uint32_t n_fill = 0; uint64_t fill_items = 0;
    DeriveTask* d_derive = nullptr; uint64_t* d_derive_prefix = nullptr; uint32_t n_derive = 0; uint64_t derive_ctas = 0;
    unsigned long long* d_long64 = nullptr;   // long-interval queue of fill_multi_kernel
    // pipelined H2D: consecutive jobs in groups; a group's inputs arrive on the copy stream and
    // `done` is recorded behind them.  Task / tile ranges are those of the group's jobs.
    struct Group {
        uint32_t job_lo = 0, job_hi = 0;
        uint32_t sc_lo = 0, sc_hi = 0, fl_lo = 0, fl_hi = 0, dv_lo = 0, dv_hi = 0, pk_lo = 0, pk_hi = 0;
        uint64_t sc_items = 0, fl_items = 0, dv_ctas = 0, pk_chunks = 0, ed_ctas = 0;
        size_t fr_lo = 0, fr_hi = 0;    // the group's F, R arrays: [fr_lo, fr_hi) of d_bits
        uint32_t tile_lo = 0, tile_hi = 0;
        cudaEvent_t done = nullptr;
    };
    // host copies of the uploaded tables: they live as long as the batch, so prepare does not
    // have to wait for their (stream-ordered) upload
    std::vector<unsigned char> h_blob;
    std::vector<EventJob> h_table;
    std::vector<Group> groups;
    bool h2d_pending = false;           // the first execute has to wait for the groups' events
    // word range of the M arrays of jobs [lo, hi) (they are laid out in job order)
    size_t jobs_m_lo(uint32_t lo, uint32_t hi) const
    {
        for (uint32_t i = lo; i < hi; i++)
            if (jobs[i].autocorr || (jobs[i].mode & MODE_MSCC)) return jobs[i].bits_m;
        return 0;
    }
    size_t jobs_m_hi(uint32_t lo, uint32_t hi) const
    {
        for (uint32_t i = hi; i > lo; i--)
            if (jobs[i - 1].autocorr || (jobs[i - 1].mode & MODE_MSCC)) return jobs[i - 1].bits_m + jobs[i - 1].array_words;
        return 0;
    }
};

// ------------------------------------------------------------------ context
extern "C" int pmb_create(pmb_ctx** out, int device_ordinal)
{
    if (!out) return PMB_ERR_ARG;
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n <= 0) {
        g_create_error = std::string("no usable CUDA device: ") +
                         (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
        return PMB_ERR_CUDA;
    }
    if (device_ordinal < 0 || device_ordinal >= n) {
        g_create_error = "device ordinal out of range";
        return PMB_ERR_ARG;
    }
    pmb_ctx* c = new pmb_ctx();
    c->device = device_ordinal;
    auto fail = [&](const char* what, cudaError_t err) {
        g_create_error = std::string(what) + ": " + cudaGetErrorString(err);
        delete c;
        return PMB_ERR_CUDA;
    };
    if ((e = cudaSetDevice(device_ordinal)) != cudaSuccess) return fail("cudaSetDevice", e);
    cudaDeviceProp prop;
    if ((e = cudaGetDeviceProperties(&prop, device_ordinal)) != cudaSuccess)
        return fail("cudaGetDeviceProperties", e);
    if (prop.major != 10) {
        g_create_error = "libpymasc_b200 is built for sm_100a only; device is sm_" +
                         std::to_string(prop.major) + std::to_string(prop.minor);
        delete c;
        return PMB_ERR_CUDA;
    }
    c->num_sms = prop.multiProcessorCount;
    if ((e = cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking)) != cudaSuccess)
        return fail("cudaStreamCreate", e);
    c->stream = c->own_stream;
    if ((e = cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking)) != cudaSuccess)
        return fail("cudaStreamCreate", e);
    if ((e = cudaStreamCreateWithFlags(&c->side_stream, cudaStreamNonBlocking)) != cudaSuccess)
        return fail("cudaStreamCreate", e);
    if ((e = cudaEventCreateWithFlags(&c->ev_alloc, cudaEventDisableTiming)) != cudaSuccess)
        return fail("cudaEventCreate", e);
    if ((e = cudaEventCreateWithFlags(&c->ev_side, cudaEventDisableTiming)) != cudaSuccess)
        return fail("cudaEventCreate", e);
    for (auto& ev : c->ev)
        if ((e = cudaEventCreate(&ev)) != cudaSuccess) return fail("cudaEventCreate", e);
    for (auto& ev : c->evk)
        if ((e = cudaEventCreate(&ev)) != cudaSuccess) return fail("cudaEventCreate", e);
    // keep freed blocks cached in the stream-ordered pool (arena behaviour)
    cudaMemPool_t pool;
    if (cudaDeviceGetDefaultMemPool(&pool, device_ordinal) == cudaSuccess) {
        uint64_t thr = UINT64_MAX;
        cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr);
    }
    *out = c;
    return PMB_OK;
}

extern "C" int pmb_device_count(void)
{
    int n = 0;
    return cudaGetDeviceCount(&n) == cudaSuccess ? n : 0;
}

extern "C" void pmb_comm_destroy(pmb_ctx* c);
extern "C" void pmb_destroy(pmb_ctx* c)
{
    if (!c) return;
    pmb_comm_destroy(c);
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    for (auto& ev : c->ev)
        if (ev) cudaEventDestroy(ev);
    for (auto& ev : c->evk)
        if (ev) cudaEventDestroy(ev);
    if (c->h_stage) cudaFreeHost(c->h_stage);
    if (c->ev_alloc) cudaEventDestroy(c->ev_alloc);
    if (c->ev_side) cudaEventDestroy(c->ev_side);
    if (c->side_stream) { cudaStreamSynchronize(c->side_stream); cudaStreamDestroy(c->side_stream); }
    if (c->copy_stream) { cudaStreamSynchronize(c->copy_stream); cudaStreamDestroy(c->copy_stream); }
    if (c->own_stream) cudaStreamDestroy(c->own_stream);
    delete c;
}

extern "C" const char* pmb_last_error(pmb_ctx* c)
{
    return c ? c->err.c_str() : g_create_error.c_str();
}

extern "C" void* pmb_stream(pmb_ctx* c) { return c ? (void*)c->stream : nullptr; }

extern "C" int pmb_set_stream(pmb_ctx* c, void* stream)
{
    if (!c) return PMB_ERR_ARG;
    CU_TRY(c, cudaSetDevice(c->device));
    CU_TRY(c, cudaStreamSynchronize(c->stream));   // nothing of ours is left on the old stream
    c->stream = stream ? (cudaStream_t)stream : c->own_stream;
    return PMB_OK;
}

extern "C" int pmb_synchronize(pmb_ctx* c)
{
    if (!c) return PMB_ERR_ARG;
    CU_TRY(c, cudaSetDevice(c->device));
    CU_TRY(c, cudaStreamSynchronize(c->stream));
    return PMB_OK;
}

extern "C" int pmb_host_alloc(void** out, uint64_t bytes)
{
    if (!out) return PMB_ERR_ARG;
    return cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocDefault) == cudaSuccess ? PMB_OK
                                                                                      : PMB_ERR_NOMEM;
}
extern "C" void pmb_host_free(void* p)
{
    if (p) cudaFreeHost(p);
}

extern "C" int pmb_set_option(pmb_ctx* c, const char* name, int64_t value)
{
    if (!c || !name) return PMB_ERR_ARG;
    if (!strcmp(name, "variant")) c->variant = (int)value;
    else if (!strcmp(name, "tile_words")) c->tile_words = (int)value;
    else if (!strcmp(name, "grid_blocks")) c->grid_blocks = (int)value;
    else if (!strcmp(name, "block_warps")) c->block_warps = (int)value;
    else if (!strcmp(name, "path")) c->path = (int)value;
    else if (!strcmp(name, "event_tile_words")) c->event_tile_words = (int)value;
    else if (!strcmp(name, "check_selftest")) c->check_selftest = (int)value;
    else if (!strcmp(name, "list_concurrent")) c->list_concurrent = (int)value;
    else if (!strcmp(name, "list_blocks_per_sm")) c->list_blocks_per_sm = (int)value;
    else if (!strcmp(name, "kernel_timing")) c->kernel_timing = value != 0;
    else if (!strcmp(name, "list_edge_items")) c->list_edge_items = (int)value;
    else if (!strcmp(name, "list_pair_cap")) c->list_pair_cap = (int)value;
    else if (!strcmp(name, "list_pairs_limit")) c->list_pairs_limit = (int)value;
    else if (!strcmp(name, "event_blocks_per_sm")) c->event_blocks_per_sm = (int)value;
    else if (!strcmp(name, "async_h2d")) c->async_h2d = value != 0;
    else if (!strcmp(name, "h2d_growth_x10")) c->h2d_growth_x10 = (int)std::max<int64_t>(10, std::min<int64_t>(value, 100));
    else if (!strcmp(name, "h2d_min_group_bytes")) c->h2d_min_group_bytes = std::max<int64_t>(0, value);
    else if (!strcmp(name, "h2d_groups")) c->h2d_groups = (int)std::max<int64_t>(1, std::min<int64_t>(value, 64));
    else { c->err = std::string("unknown option ") + name; return PMB_ERR_ARG; }
    return PMB_OK;
}

extern "C" int pmb_get_timing(pmb_ctx* c, pmb_timing* t)
{
    if (!c || !t) return PMB_ERR_ARG;
    if (c->timing.total_ms < 0.f && c->timing.launches) {   // events of the last execute
        CU_TRY(c, cudaSetDevice(c->device));
        CU_TRY(c, cudaEventSynchronize(c->ev[2]));
        CU_TRY(c, cudaEventElapsedTime(&c->timing.build_ms, c->ev[0], c->ev[1]));
        CU_TRY(c, cudaEventElapsedTime(&c->timing.window_ms, c->ev[1], c->ev[2]));
        CU_TRY(c, cudaEventElapsedTime(&c->timing.total_ms, c->ev[0], c->ev[2]));
        if (c->timing.path == 2 || (c->timing.path == 3 && c->timing_split)) {
            CU_TRY(c, cudaEventElapsedTime(&c->timing.ev_fwd_ms, c->ev[1], c->evk[0]));
            CU_TRY(c, cudaEventElapsedTime(&c->timing.ev_rev_ms, c->evk[0], c->evk[1]));
            CU_TRY(c, cudaEventElapsedTime(&c->timing.ev_edge_ms, c->evk[1], c->ev[2]));
        }
    }
    *t = c->timing;
    return PMB_OK;
}

// ------------------------------------------------------------------ batch: prepare
static uint32_t gcd_u32(uint32_t a, uint32_t b) { return b ? gcd_u32(b, a % b) : a; }

enum { A_F = 0, A_R = 1, A_M = 2, A_MP = 3, A_FM = 4 };

static bool whole_range(const JobPlan& j) { return j.lo == 0 && j.hi == (int64_t)j.nbits; }

// event-window path: array slots, in units of array_words (Er and Or are half arrays)
enum { E_F = 0, E_R = 1, E_M = 2, E_ER = 3 };
static int arrays_of(const JobPlan& j)
{
    // autocorrelation: M (whole window, the "R" role) [+ M clipped to the shard, the "F" role]
    if (j.autocorr) return whole_range(j) ? 1 : 2;
    return (j.mode & MODE_MSCC) ? 5 : 2;      // F, R [, M, Mp, FM]
}

extern "C" void pmb_batch_free(pmb_ctx* c, pmb_batch* b)
{
    if (!b) return;
    if (c) {
        cudaSetDevice(c->device);
        if (b->h2d_pending && c->copy_stream) cudaStreamSynchronize(c->copy_stream);   // never executed
        for (auto& g : b->groups)
            if (g.done) cudaEventDestroy(g.done);
        if (b->d_in) cudaFreeAsync(b->d_in, c->stream);
        if (b->d_bits_raw) cudaFreeAsync(b->d_bits_raw, c->stream);
        if (b->d_counters) cudaFreeAsync(b->d_counters, c->stream);
        if (b->d_jobs) cudaFreeAsync(b->d_jobs, c->stream);
        if (b->d_err) cudaFreeAsync(b->d_err, c->stream);
        if (b->d_long) cudaFreeAsync(b->d_long, c->stream);
        if (b->d_aux) cudaFreeAsync(b->d_aux, c->stream);
        if (b->d_evjobs) cudaFreeAsync(b->d_evjobs, c->stream);
        if (b->d_tasks) cudaFreeAsync(b->d_tasks, c->stream);
        if (b->d_long64) cudaFreeAsync(b->d_long64, c->stream);
        if (b->d_runs) cudaFreeAsync(b->d_runs, c->stream);
        if (b->d_listjobs) cudaFreeAsync(b->d_listjobs, c->stream);
        if (b->d_list_prefix) cudaFreeAsync(b->d_list_prefix, c->stream);
        if (b->d_list_acc) cudaFreeAsync(b->d_list_acc, c->stream);
        cudaStreamSynchronize(c->stream);
    }
    delete b;
}

// One shift-window launch per (mode present in the batch) x (chunk of <= 64 units).
// A unit chunk starting at unit u0 is the same kernel on moved base pointers:
// R + 4*u0 words, Mp - 4*u0 words, counters + 128*u0.
static int plan_windows(pmb_ctx* c, pmb_batch* b)
{
    const size_t n = b->jobs.size();
    const uint32_t H = b->halo_words, T = b->tile_words;
    const uint32_t UNIT_CHUNK = 64;
    std::vector<WindowJob> table;
    size_t jobs_used = 0;
    for (int mode = 1; mode <= 3; mode++) {
        std::vector<size_t> members;
        for (size_t i = 0; i < n; i++)
            if (b->jobs[i].mode == mode) members.push_back(i);
        if (members.empty()) continue;
        for (uint32_t u0 = 0; u0 < b->n_units; u0 += UNIT_CHUNK) {
            const uint32_t nu = std::min(UNIT_CHUNK, b->n_units - u0);
            table.clear();
            uint32_t tiles = 0;
            for (size_t i : members) {
                JobPlan& j = b->jobs[i];
                uint32_t* base = b->d_bits + j.bits;
                auto arr = [&](int slot) { return base + (size_t)slot * j.array_words + H; };
                WindowJob w{};
                if (j.autocorr) {   // sum_j M[j] M[j+k]: the NCC kernel on F = R = M, into the dmp slot
                    w.F = arr(whole_range(j) ? 0 : 1);
                    w.R = arr(0) + 4 * u0;
                    w.counters = b->d_counters + (i * PMB_NCOUNTERS + C_DMP) * b->stride + 128 * u0;
                } else {
                    w.F = arr(A_F);
                    w.R = arr(A_R) + 4 * u0;
                    if (j.mode & MODE_MSCC) {
                        w.FM = arr(A_FM);
                        w.M = arr(A_M);
                        w.Mp = arr(A_MP) - 4 * u0;
                    }
                    w.counters = b->d_counters + i * PMB_NCOUNTERS * b->stride + 128 * u0;
                }
                w.tile_begin = tiles;
                w.n_tiles = j.n_tiles;
                tiles += j.n_tiles;
                table.push_back(w);
            }
            if (jobs_used + table.size() > b->d_jobs_count) { c->err = "job table overflow"; return PMB_ERR_ARG; }
            WindowJob* d_tab = b->d_jobs + jobs_used;
            jobs_used += table.size();
            // pageable source: staged before the call returns, so `table` can be reused
            CU_TRY(c, cudaMemcpyAsync(d_tab, table.data(), table.size() * sizeof(WindowJob),
                                      cudaMemcpyHostToDevice, c->stream));
            WinLaunch L{};
            L.mode = mode;
            L.P.jobs = d_tab;
            L.P.n_jobs = (uint32_t)table.size();
            L.P.total_tiles = tiles;
            L.P.tile_words = T;
            L.P.n_units = nu;
            L.warps = (uint32_t)std::max(1, std::min(c->block_warps, 16));
            L.P.n_parts = L.warps / gcd_u32(nu, L.warps);
            L.P.n_shifts = b->n_shifts;
            L.P.stride = b->stride;
            L.P.halo_words = 4 * nu + 4;
            uint32_t blocks = c->grid_blocks > 0 ? (uint32_t)c->grid_blocks : (uint32_t)c->num_sms;
            L.blocks = std::max(1u, std::min(blocks, tiles));
            L.smem = window_smem_bytes(T, L.P.halo_words, nu);
            L.pairs = (uint64_t)tiles * T * nu * 128;
            b->launches.push_back(L);
        }
    }
    return PMB_OK;
}

// Event-window path: one job table for the three event kernels; tiles cover the WHOLE array of
// every job (halos included: reverse reads in the halo contribute to in-range positions).
static int plan_events(pmb_ctx* c, pmb_batch* b)
{
    const size_t n = b->jobs.size();
    const uint32_t TW = EV_TILE;
    b->ev_tile_words = TW;
    b->aux_words = n * (2 * (size_t)b->stride + 4);      // edge sums, direct lags (scatter | gather build)
    CU_TRY(c, cudaMallocAsync((void**)&b->d_aux, b->aux_words * 8, c->stream));
    CU_TRY(c, cudaMallocAsync((void**)&b->d_evjobs, n * sizeof(EventJob), c->stream));
    std::vector<EventJob>& table = b->h_table;
    table.assign(n, EventJob{});
    uint32_t tiles = 0;
    unsigned long long* adirect0 = b->d_aux + n * 2 * (size_t)b->stride;
    for (size_t i = 0; i < n; i++) {
        JobPlan& j = b->jobs[i];
        uint32_t* base = b->d_bits + j.bits;
        auto arr = [&](int slot) { return base + (size_t)slot * j.array_words; };
        EventJob e{};
        if (j.autocorr) {
            e.M = b->d_bits + j.bits_m;
            e.flags = EVF_AUTOCORR;
            e.edge_off = 0;
        } else {
            e.F = arr(E_F); e.R = arr(E_R);
            e.flags = (j.mode & MODE_NCC ? EVF_NCC : 0) | (j.mode & MODE_MSCC ? EVF_MSCC : 0);
            if (j.mode & MODE_MSCC) {
                e.M = b->d_bits + j.bits_m;
                e.Er = b->d_bits + j.bits_d; e.Or = e.Er + j.array_words / 2;
            }
            e.edge_off = (int32_t)b->read_len - 1 - (int32_t)b->max_shift;
        }
        e.counters = b->d_counters + i * PMB_NCOUNTERS * b->stride;
        // a job that shares its track reads the owner's edge sums and direct lags: finalize_dmp
        // then writes the same dmp into its own counters, and nothing is built or scanned twice
        const size_t io = j.owner >= 0 ? (size_t)j.owner : i;
        if (j.owner >= 0) e.flags |= EVF_SHARED;
        e.aux = b->d_aux + io * 2 * (size_t)b->stride;
        e.adirect = adirect0 + 2 * io;
        e.agather = adirect0 + 2 * n + 2 * io;
        e.totals = b->d_counters + n * PMB_NCOUNTERS * b->stride + 2 * i;
        const int64_t off = (int64_t)b->halo_words * 32 - j.origin_word * 32;   // chrom bit -> array bit
        e.a_lo = j.lo + off;
        e.a_hi = j.hi + off;
        e.array_words = (uint32_t)j.array_words;
        e.tile_begin = tiles;
        e.n_tiles = (uint32_t)((j.array_words + TW - 1) / TW);
        tiles += e.n_tiles;
        table[i] = e;
    }
    b->ev_total_tiles = tiles;

    // batched build: one scatter / fill / derive launch for the whole batch
    {
        const int64_t NEG = -(1ll << 40), POS = (1ll << 40);
        std::vector<ScatterTask> sc; std::vector<uint64_t> scp{0};
        std::vector<FillTask> fl; std::vector<uint64_t> flp{0};
        std::vector<DeriveTask> dv; std::vector<uint64_t> dvp{0};
        std::vector<PackTask> pk; std::vector<uint64_t> pkp{0};
        std::vector<uint64_t> edp{0};
        std::vector<uint32_t> pk_at(n + 1);
        unsigned long long* adir0 = b->d_aux + n * 2 * (size_t)b->stride;
        std::vector<uint32_t> sc_at(n + 1), fl_at(n + 1), dv_at(n + 1);
        for (size_t i = 0; i < n; i++) {
            JobPlan& j = b->jobs[i];
            sc_at[i] = (uint32_t)sc.size(); fl_at[i] = (uint32_t)fl.size(); dv_at[i] = (uint32_t)dv.size();
            pk_at[i] = (uint32_t)pk.size();
            Geometry g;
            g.origin_bit = j.origin_word * 32;
            g.halo_bits = (int64_t)b->halo_words * 32;
            g.total_bits = (int64_t)j.array_words * 32;
            uint32_t* base = b->d_bits + j.bits;
            auto arr = [&](int slot) { return base + (size_t)slot * j.array_words; };
            const bool mscc = !j.autocorr && (j.mode & MODE_MSCC);
            if (!j.autocorr) {
                if (j.a.n_fwd) { sc.push_back({arr(E_F), g, b->d_in + j.in_fwd, j.lo, j.hi, j.nbits}); scp.push_back(scp.back() + j.a.n_fwd); }
                if (j.a.n_rev) { sc.push_back({arr(E_R), g, b->d_in + j.in_rev, NEG, POS, j.nbits}); scp.push_back(scp.back() + j.a.n_rev); }
                // packed build: every F and R array is a task (an array without reads is all zeros)
                pk.push_back({arr(E_F), g, b->d_in + j.in_fwd, j.a.n_fwd, j.lo, j.hi, j.nbits, (uint32_t)j.array_words, 0u});
                pkp.push_back(pkp.back() + std::max<uint64_t>(1, (j.a.n_fwd + PACK_CHUNK - 1) / PACK_CHUNK));
                pk.push_back({arr(E_R), g, b->d_in + j.in_rev, j.a.n_rev, NEG, POS, j.nbits, (uint32_t)j.array_words, 0u});
                pkp.push_back(pkp.back() + std::max<uint64_t>(1, (j.a.n_rev + PACK_CHUNK - 1) / PACK_CHUNK));
            }
            uint32_t* M = (j.autocorr || mscc) ? b->d_bits + j.bits_m : nullptr;
            if (!M || j.owner >= 0) continue;       // a shared track is built by its owner
            if (j.a.n_intervals) {
                FillTask f{M, g, b->d_in + j.in_mb, b->d_in + j.in_me, 0, NEG, POS, j.nbits};
                f.c_lo = j.lo; f.c_hi = j.hi;
                f.p_lo = g.origin_bit - g.halo_bits; f.p_hi = f.p_lo + g.total_bits;
                f.lag0 = j.autocorr ? 0 : (long long)b->read_len - 1;
                f.lag1 = j.autocorr ? 1 : (long long)b->read_len - 2;
                f.agather = adir0 + 2 * n + 2 * i;
                fl.push_back(f);
                flp.push_back(flp.back() + j.a.n_intervals);
            }
            DeriveTask d{};
            d.M = M;
            d.Er = mscc ? b->d_bits + j.bits_d : nullptr;
            d.Or = mscc ? d.Er + j.array_words / 2 : nullptr;
            d.adirect = adir0 + 2 * i;
            const int64_t off = g.halo_bits - g.origin_bit;
            d.a_lo = j.lo + off; d.a_hi = j.hi + off;
            d.lag0 = j.autocorr ? 0 : (long long)b->read_len - 1;
            d.lag1 = j.autocorr ? 1 : (long long)b->read_len - 2;
            d.array_words = (uint32_t)j.array_words;
            d.n_intervals = (uint32_t)j.a.n_intervals;
            d.begins = b->d_in + j.in_mb;
            d.ends = b->d_in + j.in_me;
            d.p_off = g.origin_bit - g.halo_bits;
            d.aux = b->d_aux + i * 2 * (size_t)b->stride;
            d.edge_off = j.autocorr ? 0 : (int)b->read_len - 1 - (int)b->max_shift;
            dv.push_back(d);
            edp.push_back(edp.back() + (j.a.n_intervals + EDGE_INTERVALS_PER_CTA - 1) / EDGE_INTERVALS_PER_CTA);
            dvp.push_back(dvp.back() + (j.array_words / 4 + DERIVE_QUADS_PER_CTA - 1) / DERIVE_QUADS_PER_CTA);
        }
        sc_at[n] = (uint32_t)sc.size(); fl_at[n] = (uint32_t)fl.size(); dv_at[n] = (uint32_t)dv.size();
        pk_at[n] = (uint32_t)pk.size();
        for (auto& gr : b->groups) {
            gr.sc_lo = sc_at[gr.job_lo]; gr.sc_hi = sc_at[gr.job_hi]; gr.sc_items = scp[gr.sc_hi] - scp[gr.sc_lo];
            gr.fl_lo = fl_at[gr.job_lo]; gr.fl_hi = fl_at[gr.job_hi]; gr.fl_items = flp[gr.fl_hi] - flp[gr.fl_lo];
            gr.dv_lo = dv_at[gr.job_lo]; gr.dv_hi = dv_at[gr.job_hi]; gr.dv_ctas = dvp[gr.dv_hi] - dvp[gr.dv_lo];
            gr.pk_lo = pk_at[gr.job_lo]; gr.pk_hi = pk_at[gr.job_hi]; gr.pk_chunks = pkp[gr.pk_hi] - pkp[gr.pk_lo];
            gr.ed_ctas = edp[gr.dv_hi] - edp[gr.dv_lo];
            gr.fr_lo = gr.fr_hi = 0;
            for (uint32_t q = gr.job_lo; q < gr.job_hi; q++)
                if (!b->jobs[q].autocorr) {
                    if (gr.fr_hi == 0) gr.fr_lo = b->jobs[q].bits;
                    gr.fr_hi = b->jobs[q].bits + 2 * b->jobs[q].array_words;
                }
            for (uint32_t q = gr.job_lo; q < gr.job_hi; q++) table[q].gate_index = (uint32_t)(&gr - b->groups.data());
            gr.tile_lo = table[gr.job_lo].tile_begin;
            gr.tile_hi = table[gr.job_hi - 1].tile_begin + table[gr.job_hi - 1].n_tiles;
        }
        b->n_scatter = (uint32_t)sc.size(); b->scatter_items = scp.back();
        b->n_fill = (uint32_t)fl.size(); b->fill_items = flp.back();
        b->n_derive = (uint32_t)dv.size(); b->derive_ctas = dvp.back();
        b->n_pack = (uint32_t)pk.size(); b->pack_chunks = pkp.back();
        b->edge_ctas = edp.back();
        auto bytes8 = [](size_t x) { return (x + 15) & ~(size_t)15; };
        const size_t o_sc = 0, o_scp = o_sc + bytes8(sc.size() * sizeof(ScatterTask));
        const size_t o_fl = o_scp + bytes8(scp.size() * 8), o_flp = o_fl + bytes8(fl.size() * sizeof(FillTask));
        const size_t o_dv = o_flp + bytes8(flp.size() * 8), o_dvp = o_dv + bytes8(dv.size() * sizeof(DeriveTask));
        const size_t o_pk = o_dvp + bytes8(dvp.size() * 8), o_pkp = o_pk + bytes8(pk.size() * sizeof(PackTask));
        const size_t o_edp = o_pkp + bytes8(pkp.size() * 8);
        const size_t total = o_edp + bytes8(edp.size() * 8);
        std::vector<unsigned char>& blob = b->h_blob;
        blob.assign(total, 0);
        if (!sc.empty()) memcpy(blob.data() + o_sc, sc.data(), sc.size() * sizeof(ScatterTask));
        memcpy(blob.data() + o_scp, scp.data(), scp.size() * 8);
        if (!fl.empty()) memcpy(blob.data() + o_fl, fl.data(), fl.size() * sizeof(FillTask));
        memcpy(blob.data() + o_flp, flp.data(), flp.size() * 8);
        if (!dv.empty()) memcpy(blob.data() + o_dv, dv.data(), dv.size() * sizeof(DeriveTask));
        memcpy(blob.data() + o_dvp, dvp.data(), dvp.size() * 8);
        if (!pk.empty()) memcpy(blob.data() + o_pk, pk.data(), pk.size() * sizeof(PackTask));
        memcpy(blob.data() + o_pkp, pkp.data(), pkp.size() * 8);
        memcpy(blob.data() + o_edp, edp.data(), edp.size() * 8);
        CU_TRY(c, cudaMallocAsync(&b->d_tasks, total, c->stream));
        CU_TRY(c, cudaMemcpyAsync(b->d_tasks, blob.data(), total, cudaMemcpyHostToDevice, c->stream));
        unsigned char* d = (unsigned char*)b->d_tasks;
        b->d_scatter = (ScatterTask*)(d + o_sc); b->d_scatter_prefix = (uint64_t*)(d + o_scp);
        b->d_fill = (FillTask*)(d + o_fl); b->d_fill_prefix = (uint64_t*)(d + o_flp);
        b->d_derive = (DeriveTask*)(d + o_dv); b->d_derive_prefix = (uint64_t*)(d + o_dvp);
        b->d_pack = (PackTask*)(d + o_pk); b->d_pack_prefix = (uint64_t*)(d + o_pkp);
        b->d_edge_prefix = (uint64_t*)(d + o_edp);
        b->runs_cap = b->m_begin / PACK_LONG_GAP + pk.size() + 16;
        CU_TRY(c, cudaMallocAsync((void**)&b->d_runs, 2 * b->runs_cap * sizeof(ZeroRun), c->stream));
        CU_TRY(c, cudaMallocAsync((void**)&b->d_long64, 2 * (b->fill_items + 1) * 8, c->stream));   // main | side
    }
    CU_TRY(c, cudaMemcpyAsync(b->d_evjobs, table.data(), n * sizeof(EventJob), cudaMemcpyHostToDevice,
                              c->stream));
    return PMB_OK;
}

// Geometry of the bit arrays of the dense / event path (b->path says which): tile size, halos,
// where every job's arrays start.  The packed inputs are laid out by prepare_impl.
static int layout_arrays(pmb_ctx* c, pmb_batch* b)
{
    const uint32_t blocks = c->grid_blocks > 0 ? (uint32_t)c->grid_blocks : (uint32_t)c->num_sms;
    b->halo_words = 4 * b->n_units + 4;
    uint32_t T = 1024;
    if (b->path == 2) {
        T = 64;   // body granule only; the event kernels tile the whole array themselves
        // windows reach 2*max_shift + read_len beyond the counted range (the stride-2 window of
        // the reverse reads), plus the idle-lane / upper-neighbour words
        b->halo_words = ((2 * b->max_shift + b->read_len + 192 + 31) / 32 + 3) & ~3u;
    } else {
        if (c->tile_words > 0) T = (uint32_t)c->tile_words;
        else while (T > 64 && b->total_words / T < 8ull * blocks) T >>= 1;
        if (T < 64 || (T & 63)) { c->err = "tile_words must be a multiple of 64"; return PMB_ERR_ARG; }
    }
    b->tile_words = T;
    size_t bits_words = 0;
    for (auto& j : b->jobs) {
        uint64_t body = (uint64_t)(((j.hi + 31) >> 5) - j.origin_word);
        j.n_tiles = (uint32_t)((body + T - 1) / T);
        if (j.n_tiles == 0) j.n_tiles = 1;
        j.body_words = (uint64_t)j.n_tiles * T;
        j.array_words = 2 * (size_t)b->halo_words + j.body_words;
        j.bits = bits_words;
        if (b->path == 2) {     // F, R of all jobs first (one memset), then all M, then the derived arrays
            if (!j.autocorr) bits_words += 2 * j.array_words;
        } else {
            bits_words += (size_t)arrays_of(j) * j.array_words;
        }
    }
    if (b->path == 2) {
        b->m_begin = bits_words;
        for (auto& j : b->jobs)
            if (j.owner >= 0) j.bits_m = b->jobs[j.owner].bits_m;
            else if (j.autocorr || (j.mode & MODE_MSCC)) { j.bits_m = bits_words; bits_words += j.array_words; }
        b->m_end = bits_words;
        for (auto& j : b->jobs)
            if (j.owner >= 0) j.bits_d = b->jobs[j.owner].bits_d;
            else if (!j.autocorr && (j.mode & MODE_MSCC)) { j.bits_d = bits_words; bits_words += j.array_words; }
        // M is gathered from the sorted intervals (the scatter kernels stay as the fallback)
        b->gather_k = 0;
        b->gather_m = b->read_len >= 2;   // rl 1: the second lag is negative
        for (auto& j : b->jobs)
            if (j.a.n_intervals >= (1ull << 31)) b->gather_m = false;
        if (getenv("PMB_NO_GATHER")) b->gather_m = false;     // debug aid: always scatter M
        b->edge_list_ok = true;
        for (auto& j : b->jobs)
            if ((uint64_t)j.array_words * 32 + 32ull * b->stride >= (1ull << 32)) b->edge_list_ok = false;
        b->pack_reads = getenv("PMB_NO_PACK") == nullptr;      // debug aid: always scatter F, R
        b->zero_words = b->gather_m ? b->m_begin : b->m_end;
    }
    b->bits_words = bits_words;
    return PMB_OK;
}

// Device allocations and job / task tables of the dense or event path (after layout_arrays).
static int plan_arrays(pmb_ctx* c, pmb_batch* b)
{
    const size_t n = b->jobs.size();
    CU_TRY(c, cudaMallocAsync((void**)&b->d_bits_raw,
                              (std::max<size_t>(b->bits_words, 4) + 2 * (size_t)EV_PAD_WORDS) * 4, c->stream));
    b->d_bits = b->d_bits_raw + EV_PAD_WORDS;
    // the pads are only ever read by TMA region overhang (never used); zero them once
    CU_TRY(c, cudaMemsetAsync(b->d_bits_raw, 0, (size_t)EV_PAD_WORDS * 4, c->stream));
    CU_TRY(c, cudaMemsetAsync(b->d_bits + std::max<size_t>(b->bits_words, 4), 0, (size_t)EV_PAD_WORDS * 4, c->stream));
    uint64_t max_iv = 0;
    for (auto& j : b->jobs) max_iv = std::max<uint64_t>(max_iv, j.a.n_intervals);
    b->long_cap = (size_t)max_iv + 4;
    CU_TRY(c, cudaMallocAsync((void**)&b->d_long, b->long_cap * 4, c->stream));
    // window job tables: one per unit chunk (filled in execute)
    b->d_jobs_count = (size_t)n * ((b->n_units + 63) / 64 + 1) * 3;
    CU_TRY(c, cudaMallocAsync((void**)&b->d_jobs, b->d_jobs_count * sizeof(WindowJob), c->stream));
    const int rc = b->path == 2 ? plan_events(c, b) : plan_windows(c, b);
    if (rc == PMB_OK) b->arrays_planned = true;
    return rc;
}

// List path: job table + the three entry prefixes (forward reads | reverse reads | intervals of
// the jobs that own a track), accumulators.  No bit arrays.
static int plan_lists(pmb_ctx* c, pmb_batch* b)
{
    const size_t n = b->jobs.size();
    const size_t stride = b->stride;
    // two copies of [n][LA_N][stride] + [n][2] totals | [n][2][stride] run-edge products + [n][2] lags
    const size_t job_acc = LA_N * stride + 2, job_aux = 2 * stride + 2;
    b->list_acc_copy = n * (job_acc + job_aux);
    b->list_acc_words = 2 * b->list_acc_copy;
    CU_TRY(c, cudaMallocAsync((void**)&b->d_list_acc, b->list_acc_words * 8, c->stream));
    // cleared ONCE: from then on the finalize of every execute clears the copy the next one uses.
    // The outputs are written, not added to -- but a job outside this rank's range is never written.
    CU_TRY(c, cudaMemsetAsync(b->d_list_acc, 0, b->list_acc_words * 8, c->stream));
    CU_TRY(c, cudaMemsetAsync(b->d_counters, 0, (b->counter_words + 2) * 8, c->stream));
    CU_TRY(c, cudaMemsetAsync(b->d_err, 0, 4 * (4 + 128), c->stream));
    CU_TRY(c, cudaMallocAsync((void**)&b->d_listjobs, n * sizeof(ListJob), c->stream));
    CU_TRY(c, cudaMallocAsync((void**)&b->d_list_prefix, 3 * (n + 1) * 8, c->stream));
    unsigned long long* acc0 = b->d_list_acc;
    unsigned long long* aux0 = acc0 + n * job_acc;
    std::vector<ListJob>& table = b->h_listjobs;
    std::vector<unsigned long long>& pre = b->h_list_prefix;
    table.assign(n, ListJob{});
    pre.assign(3 * (n + 1), 0);
    for (size_t i = 0; i < n; i++) {
        const JobPlan& j = b->jobs[i];
        ListJob e{};
        const bool mscc = !j.autocorr && (j.mode & MODE_MSCC);
        if (!j.autocorr) {
            e.fwd = b->d_in + j.in_fwd; e.n_fwd = (uint32_t)j.a.n_fwd;
            e.rev = b->d_in + j.in_rev; e.n_rev = (uint32_t)j.a.n_rev;
            e.flags = (j.mode & MODE_NCC ? EVF_NCC : 0) | (mscc ? EVF_MSCC : 0);
        } else {
            e.flags = EVF_AUTOCORR;
        }
        if (mscc || j.autocorr) {
            e.begins = b->d_in + j.in_mb; e.ends = b->d_in + j.in_me; e.n_iv = (uint32_t)j.a.n_intervals;
        }
        const size_t io = j.owner >= 0 ? (size_t)j.owner : i;
        if (j.owner >= 0) e.flags |= EVF_SHARED;
        e.counters = b->d_counters + i * PMB_NCOUNTERS * stride;
        e.totals = b->d_counters + n * PMB_NCOUNTERS * stride + 2 * i;
        e.acc = acc0 + i * job_acc;
        e.aux = aux0 + io * job_aux;
        e.lo = j.lo; e.hi = j.hi; e.nbits = j.nbits;
        table[i] = e;
        pre[i + 1] = pre[i] + e.n_fwd;
        pre[(n + 1) + i + 1] = pre[(n + 1) + i] + e.n_rev;
        pre[2 * (n + 1) + i + 1] = pre[2 * (n + 1) + i] + ((mscc || j.autocorr) && j.owner < 0 ? e.n_iv : 0);
    }
    CU_TRY(c, cudaMemcpyAsync(b->d_listjobs, table.data(), n * sizeof(ListJob), cudaMemcpyHostToDevice, c->stream));
    CU_TRY(c, cudaMemcpyAsync(b->d_list_prefix, pre.data(), pre.size() * 8, cudaMemcpyHostToDevice, c->stream));
    return PMB_OK;
}

static int prepare_impl(pmb_ctx* c, const pmb_chrom_args* a, uint32_t n, pmb_batch** out, bool async)
{
    if (!c || !a || !out || n == 0) return PMB_ERR_ARG;
    *out = nullptr;
    CU_TRY(c, cudaSetDevice(c->device));
    pmb_batch* b = new pmb_batch();
    struct Guard {   // frees the batch on any early return
        pmb_ctx* c; pmb_batch* b; bool armed = true;
        ~Guard() { if (armed) pmb_batch_free(c, b); }
    } guard{c, b};

    b->read_len = a[0].read_len;
    b->max_shift = a[0].max_shift;
    b->n_shifts = b->max_shift + 1;
    b->n_units = (b->n_shifts + 127) / 128;
    b->stride = 128 * b->n_units;
    b->halo_words = 4 * b->n_units + 4;
    if (b->read_len == 0) { c->err = "read_len must be >= 1"; return PMB_ERR_ARG; }

    // pass 1: ranges and the total amount of work, to pick the tile size
    uint64_t total_words = 0;
    b->jobs.resize(n);
    for (uint32_t i = 0; i < n; i++) {
        JobPlan& j = b->jobs[i];
        j.a = a[i];
        if (a[i].read_len != b->read_len || a[i].max_shift != b->max_shift) {
            c->err = "all chromosomes of a batch must share read_len and max_shift";
            return PMB_ERR_ARG;
        }
        j.nbits = a[i].chrom_len + a[i].read_len + a[i].max_shift + PMB_EXTRA_ALLOCATE_SIZE;
        if (j.nbits >= (1ull << 32)) {
            c->err = "chromosome too long for 32-bit positions";
            return PMB_ERR_ARG;
        }
        j.autocorr = (a[i].flags & PMB_FLAG_AUTOCORR) != 0;
        j.mode = (a[i].flags & PMB_FLAG_NCC ? MODE_NCC : 0) | (a[i].flags & PMB_FLAG_MSCC ? MODE_MSCC : 0);
        if (j.autocorr) j.mode = MODE_NCC;
        if (j.mode == 0) { c->err = "flags select nothing to compute"; return PMB_ERR_ARG; }
        if ((j.mode & MODE_MSCC) && !(a[i].flags & PMB_FLAG_SAME_TRACK) && a[i].n_intervals &&
            (!a[i].map_begin || !a[i].map_end)) {
            c->err = "interval arrays missing"; return PMB_ERR_ARG;
        }
        if ((a[i].n_fwd && !a[i].fwd_pos) || (a[i].n_rev && !a[i].rev_pos)) {
            c->err = "position arrays missing"; return PMB_ERR_ARG;
        }
        j.lo = (int64_t)a[i].range_lo;
        j.hi = a[i].range_hi ? (int64_t)a[i].range_hi : (int64_t)j.nbits;
        if (j.hi > (int64_t)j.nbits) j.hi = (int64_t)j.nbits;
        if (j.lo > j.hi) j.lo = j.hi;
        if (a[i].flags & PMB_FLAG_SAME_TRACK) {
            // same chromosome, range and intervals as the previous job (another read set over the
            // same mappability track): M, Er/Or and dmp are built once, by the first job of the run
            const JobPlan* p = i ? &b->jobs[i - 1] : nullptr;
            if (!p || j.autocorr || !(j.mode & MODE_MSCC) || p->autocorr || !(p->mode & MODE_MSCC) ||
                p->nbits != j.nbits || p->lo != j.lo || p->hi != j.hi) {
                c->err = "PMB_FLAG_SAME_TRACK needs a preceding MSCC job with the same chrom_len and range";
                return PMB_ERR_ARG;
            }
            j.owner = p->owner >= 0 ? p->owner : (int)i - 1;
            j.a.n_intervals = b->jobs[j.owner].a.n_intervals;
            j.a.map_begin = j.a.map_end = nullptr;
        }
        j.origin_word = (j.lo >> 5) & ~(int64_t)3;
        total_words += (uint64_t)(((j.hi + 31) >> 5) - j.origin_word);
    }
    // Which algorithm?  All three give identical counts.
    //   lists  (3): cost ~ reads + run edges + PAIRS of entries within max_shift of each other
    //   events (2): cost ~ (reads + run edges) x shifts/32 + a scan of the arrays
    //   dense  (1): cost ~ words x shifts
    // auto: lists while a forward read expects fewer than ~64 reverse reads per 1024 shifts
    // (measured crossover ~75: profiles/r02_config4_chr1_paths.jsonl),
    // else events unless the inputs hold more than 4 events per 32-bit word.
    uint64_t total_events = 0;
    double expect_pairs = 0, total_fwd = 0;
    b->list_ok = b->read_len >= 2 && list_smem_bytes(b->max_shift, list_fwd_hists(true)) <= 200 * 1024;
    for (uint32_t i = 0; i < n; i++) {
        total_events += a[i].n_fwd + a[i].n_rev + (b->jobs[i].owner < 0 ? 2 * a[i].n_intervals : 0);
        b->any_ncc |= (a[i].flags & PMB_FLAG_NCC) && !(a[i].flags & PMB_FLAG_AUTOCORR);
        b->any_mscc |= (a[i].flags & PMB_FLAG_MSCC) && !(a[i].flags & PMB_FLAG_AUTOCORR);
        b->any_autocorr |= (a[i].flags & PMB_FLAG_AUTOCORR) != 0;
        if (a[i].n_fwd >= (1ull << 31) || a[i].n_rev >= (1ull << 31) || a[i].n_intervals >= (1ull << 31) ||
            b->jobs[i].nbits + 2ull * (b->read_len + b->max_shift) + 64 >= (1ull << 32))   // (distances modulo 2^32 in the list kernels)
            b->list_ok = false;
        if (!(a[i].flags & PMB_FLAG_AUTOCORR)) {
            expect_pairs += (double)a[i].n_fwd * (double)a[i].n_rev * (double)b->n_shifts / (double)b->jobs[i].nbits;
            total_fwd += (double)a[i].n_fwd;
        }
    }
    if (getenv("PMB_NO_LISTS")) b->list_ok = false;         // debug aid
    const double pairs_limit = (c->list_pairs_limit > 0 ? c->list_pairs_limit : 64) * ((b->n_shifts + 1023) / 1024);
    if (c->path == 1 || c->path == 2) b->path = c->path;
    else if (c->path == 3) {
        if (!b->list_ok) { c->err = "the list path needs read_len >= 2, max_shift <= 17000 and fewer than 2^31 entries per list"; return PMB_ERR_ARG; }
        b->path = 3;
    } else if (b->list_ok && expect_pairs <= pairs_limit * (total_fwd + 1.0)) b->path = 3;
    else b->path = total_events <= 4 * total_words ? 2 : 1;
    b->total_words = total_words;

    // pass 2: layout of the packed inputs on the device (the bit arrays: layout_arrays)
    size_t in_words = 0;
    for (uint32_t i = 0; i < n; i++) {
        JobPlan& j = b->jobs[i];
        auto take = [&](uint64_t cnt) { size_t o = in_words; in_words += (cnt + 3) & ~3ull; return o; };
        if (!j.autocorr) { j.in_fwd = take(j.a.n_fwd); j.in_rev = take(j.a.n_rev); }
        if (j.owner >= 0) { j.in_mb = b->jobs[j.owner].in_mb; j.in_me = b->jobs[j.owner].in_me; }
        else if ((j.mode & MODE_MSCC) || j.autocorr) { j.in_mb = take(j.a.n_intervals); j.in_me = take(j.a.n_intervals); }
    }
    b->in_words = in_words;
    b->counter_words = (size_t)n * PMB_NCOUNTERS * b->stride + (size_t)n * 2;
    if (b->path != 3) {
        int rc = layout_arrays(c, b);
        if (rc != PMB_OK) return rc;
    }

    CU_TRY(c, cudaMallocAsync((void**)&b->d_in, std::max<size_t>(in_words, 4) * 4, c->stream));
    // H2D of the packed inputs.  Synchronous mode: on the compute stream, complete when prepare
    // returns.  Asynchronous mode: on the copy stream in groups of consecutive jobs of about equal
    // bytes, an event behind each group; the first execute consumes group g while g+1 is in flight.
    cudaStream_t up_stream = c->stream;
    {
        std::vector<size_t> job_words(n);
        for (uint32_t i = 0; i < n; i++) {
            const JobPlan& j = b->jobs[i];
            job_words[i] = (j.autocorr ? 0 : j.a.n_fwd + j.a.n_rev) +
                           ((((j.mode & MODE_MSCC) || j.autocorr) && j.owner < 0) ? 2 * j.a.n_intervals : 0);
        }
        size_t total = 0, acc = 0;
        for (size_t w : job_words) total += w;
        // a group is only worth its extra launches when its copy takes a few hundred microseconds:
        // at most "h2d_groups" groups, and none smaller than ~16 MB on average
        uint32_t want = 1;
        if (async && b->path != 1) {
            // (list path: its kernels run ~3x faster than the copy, so what is exposed is the LAST
            // group's kernels after the copy has ended: twice the groups at half the size)
            const size_t max_groups = (size_t)c->h2d_groups * (b->path == 3 ? 2 : 1);
            const size_t min_bytes = (size_t)c->h2d_min_group_bytes / (b->path == 3 ? 2 : 1);
            want = c->h2d_min_group_bytes > 0
                       ? (uint32_t)std::min<size_t>(max_groups, std::max<size_t>(1, total * 4 / std::max<size_t>(min_bytes, 1)))
                       : (uint32_t)max_groups;
            want = std::min<uint32_t>(want, 64);
        }
        // Group sizes grow geometrically: the first group is small so the kernels start early,
        // and each later group may be larger because computing on a group takes ~2x as long
        // as copying it (so the copy of group g+1 hides behind the kernels of group g).
        // (The list kernels are ~10x faster than the copy: equal groups, the last one's kernels
        // are all that is left after the copy ends.)
        const double growth = b->path == 3 ? 1.0 : c->h2d_growth_x10 / 10.0;
        std::vector<double> cum(want, 1.0);
        {
            double sum = 0, w = 1;
            for (uint32_t g = 0; g < want; g++, w *= growth) sum += w;
            double run = 0; w = 1;
            for (uint32_t g = 0; g < want; g++, w *= growth) { run += w; cum[g] = run / sum; }
        }
        pmb_batch::Group g;
        for (uint32_t i = 0; i < n; i++) {
            acc += job_words[i];
            const uint32_t made = (uint32_t)b->groups.size();
            // close the group once its share of the bytes is in (the last group takes the rest)
            // (never in front of a job that shares the track of this group's jobs)
            if (i + 1 == n || (made + 1 < want && (double)acc >= cum[made] * (double)total &&
                               b->jobs[i + 1].owner < 0)) {
                g.job_hi = i + 1;
                b->groups.push_back(g);
                g = pmb_batch::Group();
                g.job_lo = i + 1;
            }
        }
        // Everything else the batch needs on the device -- arrays, counters, and the job / task
        // tables of plan_events -- is allocated and uploaded BEFORE the input copy is queued: the
        // table uploads are small H2D copies on the compute stream, and behind 90 MB .. 1.2 GB of
        // inputs on the same copy engine they held the first kernel back by the whole input copy
        // (measured: 16 read sets, execute started 22 ms late; one genome, ~1 ms).
        CU_TRY(c, cudaMallocAsync((void**)&b->d_counters, (b->counter_words + 2) * 8, c->stream));   // + the all-reduce's status words
        CU_TRY(c, cudaMallocAsync((void**)&b->d_err, 4 * (4 + 128), c->stream));
        {
            int rc = b->path == 3 ? plan_lists(c, b) : plan_arrays(c, b);
            if (rc != PMB_OK) return rc;
        }
        if (async) {
            for (auto& gr : b->groups)
                CU_TRY(c, cudaEventCreateWithFlags(&gr.done, cudaEventDisableTiming));
            CU_TRY(c, cudaEventRecord(c->ev_alloc, c->stream));        // d_in exists from here on
            CU_TRY(c, cudaStreamWaitEvent(c->copy_stream, c->ev_alloc, 0));
            up_stream = c->copy_stream;
            b->h2d_pending = true;
        }
    }
    // One cudaMemcpyAsync per run of arrays that lie back to back in host memory with the same
    // 16-byte padding as on the device (a host that packs its position arrays into one pinned
    // buffer gets one copy per group instead of four per chromosome; ~1.5 us of launch each).
    for (auto& gr : b->groups) {
        const uint32_t* run_src = nullptr; size_t run_dst = 0, run_words = 0;
        auto flush = [&]() -> cudaError_t {
            cudaError_t e = cudaSuccess;
            if (run_words)
                e = cudaMemcpyAsync(b->d_in + run_dst, run_src, run_words * 4, cudaMemcpyHostToDevice, up_stream);
            run_words = 0;
            return e;
        };
        auto up = [&](size_t off, const uint32_t* src, uint64_t cnt) -> cudaError_t {
            if (!cnt) return cudaSuccess;
            const size_t padded = (run_words + 3) & ~(size_t)3;
            if (run_words && src == run_src + padded && off == run_dst + padded) {
                run_words = padded + cnt;       // the pad words in between are the host's own padding
                return cudaSuccess;
            }
            cudaError_t e = flush();
            run_src = src; run_dst = off; run_words = cnt;
            return e;
        };
        for (uint32_t ji = gr.job_lo; ji < gr.job_hi; ji++) {
            JobPlan& j = b->jobs[ji];
            if (!j.autocorr) {
                CU_TRY(c, up(j.in_fwd, j.a.fwd_pos, j.a.n_fwd));
                CU_TRY(c, up(j.in_rev, j.a.rev_pos, j.a.n_rev));
            }
            if (((j.mode & MODE_MSCC) || j.autocorr) && j.owner < 0) {
                CU_TRY(c, up(j.in_mb, j.a.map_begin, j.a.n_intervals));
                CU_TRY(c, up(j.in_me, j.a.map_end, j.a.n_intervals));
            }
            j.a.fwd_pos = j.a.rev_pos = j.a.map_begin = j.a.map_end = nullptr;   // not read again by the host
        }
        CU_TRY(c, flush());
        if (async) CU_TRY(c, cudaEventRecord(gr.done, up_stream));
    }
    // synchronous mode: host arrays may be released by the caller; asynchronous mode: they are
    // still being read by the copy engine until the first execute of this batch has completed
    if (!async || b->path == 1) CU_TRY(c, cudaStreamSynchronize(c->stream));
    guard.armed = false;
    *out = b;
    return PMB_OK;
}

extern "C" int pmb_batch_prepare(pmb_ctx* c, const pmb_chrom_args* a, uint32_t n, pmb_batch** out)
{
    return guarded(c, [&]() -> int { return prepare_impl(c, a, n, out, c && c->async_h2d); });
}

// ------------------------------------------------------------------ batch: execute
template <int MODE>
static cudaError_t launch_window(pmb_ctx* c, int variant, const WindowParams& P, dim3 grid, dim3 block,
                                 size_t smem)
{
    // variant -> carry-save level of (ncc, dmp, mfwd, mrev, mscc)
#define PMB_LAUNCH(LV)                                                                            \
    do {                                                                                          \
        auto k = shift_window_kernel<MODE, LV>;                                                   \
        cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize,      \
                                             (int)smem);                                          \
        if (e != cudaSuccess) return e;                                                           \
        k<<<grid, block, smem, c->stream>>>(P);                                                   \
        return cudaGetLastError();                                                                \
    } while (0)
    using V0 = Levels<0, 0, 0, 0, 0>;   // plain popc everywhere (POPC-pipe bound)
    using V1 = Levels<1, 1, 1, 1, 1>;   // pair carry-save everywhere
    using V2 = Levels<2, 2, 2, 2, 2>;   // quad carry-save everywhere
    using V3 = Levels<1, 0, 1, 1, 1>;   // mixes, to balance the ALU and POPC pipes
    using V4 = Levels<1, 0, 0, 1, 1>;
    using V5 = Levels<0, 1, 1, 0, 0>;
    switch (variant) {
        case 0: PMB_LAUNCH(V0);
        case 2: PMB_LAUNCH(V2);
        case 3: PMB_LAUNCH(V3);
        case 4: PMB_LAUNCH(V4);
        case 5: PMB_LAUNCH(V5);
        case 1:
        default: PMB_LAUNCH(V1);
    }
#undef PMB_LAUNCH
}

// ------------------------------------------------------------------ batch: execute (list path)
static void list_attributes(pmb_ctx* c)
{
    // once per context (the attribute is per device, and a process may drive several): the
    // kernels' dynamic shared memory exceeds 48 KB for max_shift > ~3000
    if (c->list_attr_done) return;
    const int cap = 200 * 1024;
    cudaFuncSetAttribute(list_fwd_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, cap);
    cudaFuncSetAttribute(list_fwd_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, cap);
    cudaFuncSetAttribute(list_fwd_kernel<false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, cap);
    cudaFuncSetAttribute(list_rev_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, cap);
    cudaFuncSetAttribute(list_rev_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, cap);
    cudaFuncSetAttribute(list_edge_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, cap);
    c->list_attr_done = true;
}

static int execute_lists(pmb_ctx* c, pmb_batch* b)
{
    cudaStream_t st = c->stream;
    const size_t n = b->jobs.size();
    uint32_t launches = 0;
    list_attributes(c);
    const bool split = c->kernel_timing != 0;
    CU_TRY(c, cudaEventRecord(c->ev[0], st));
    CU_TRY(c, cudaEventRecord(c->ev[1], st));     // (no build stage: nothing is zeroed or built per execute)
    const uint32_t parity = b->list_execs & 1u;
    b->list_execs++;
    const bool grouped = b->h2d_pending && b->groups.size() > 1;
    const size_t n_pass = grouped ? b->groups.size() : 1;
    if (b->h2d_pending && !grouped)
        for (auto& gr : b->groups) CU_TRY(c, cudaStreamWaitEvent(st, gr.done, 0));
    ListParams P{};
    P.jobs = b->d_listjobs;
    P.max_shift = b->max_shift;
    P.read_len = b->read_len;
    P.stride = b->stride;
    P.pair_cap = c->list_pair_cap > 0 ? (uint32_t)c->list_pair_cap : LIST_PAIR_CAP;
    P.flags = b->d_err + 2 * parity;
    P.acc_off = parity ? b->list_acc_copy : 0;
    P.acc_other = parity ? 0 : b->list_acc_copy;
    P.flags_other = b->d_err + 2 * (parity ^ 1u);
    P.status_tail = b->d_counters + b->counter_words;
    const uint32_t bps = c->list_blocks_per_sm > 0 ? (uint32_t)c->list_blocks_per_sm : (uint32_t)LIST_MIN_BLOCKS;
    const unsigned long long* pre = b->h_list_prefix.data();
    // CTAs of at least 2048 reads -- but 512 intervals: a thread of the run-edge kernel spends its time
    // in dependent loads, so a short interval list (a rank's share) is better spread over all SMs
    const unsigned long long edge_items = c->list_edge_items > 0 ? (unsigned long long)c->list_edge_items : 512ull;
    auto grid_for = [&](unsigned long long items, unsigned long long per_cta) {
        return (unsigned)std::max<unsigned long long>(1, std::min<unsigned long long>((items + per_cta - 1) / per_cta,
                                                                                      (unsigned long long)c->num_sms * bps));
    };
    const size_t sm_f = list_smem_bytes(b->max_shift, list_fwd_hists(b->any_mscc)), sm_2 = list_smem_bytes(b->max_shift, 2);
    // The three list kernels do not depend on each other.  On long lists each fills the GPU by
    // itself and they run back to back (clean per-kernel times); on short ones -- a rank's share
    // of a genome cut into 4 or 8 ranges -- a launch is mostly ramp and tail, so the reverse and
    // edge kernels run beside the forward kernel on the side stream.
    const unsigned long long all_entries = pre[n] + pre[(n + 1) + n] + pre[2 * (n + 1) + n];
    const bool concurrent = !grouped && (c->list_concurrent < 0 ? all_entries < 8000000ull : c->list_concurrent != 0);
    if (concurrent) {
        CU_TRY(c, cudaEventRecord(c->ev_side, st));
        CU_TRY(c, cudaStreamWaitEvent(c->side_stream, c->ev_side, 0));
    }
    for (size_t pass = 0; pass < n_pass; pass++) {
        const uint32_t jl = grouped ? b->groups[pass].job_lo : 0, jh = grouped ? b->groups[pass].job_hi : (uint32_t)n;
        if (grouped) CU_TRY(c, cudaStreamWaitEvent(st, b->groups[pass].done, 0));
        for (int kind = 0; kind < 3; kind++) {      // forward reads | reverse reads | run edges
            const unsigned long long* pk = pre + kind * (n + 1);
            const unsigned long long items = pk[jh] - pk[jl];
            cudaStream_t ks = (concurrent && kind > 0) ? c->side_stream : st;
            if (items) {
                P.prefix = b->d_list_prefix + kind * (n + 1);
                P.job_lo = jl; P.job_hi = jh;
                const unsigned grid = grid_for(items, kind == 2 ? edge_items : 2048ull);
                if (kind == 0) {
                    if (b->any_ncc && b->any_mscc) list_fwd_kernel<true, true><<<grid, LIST_THREADS, sm_f, ks>>>(P);
                    else if (b->any_ncc) list_fwd_kernel<true, false><<<grid, LIST_THREADS, sm_f, ks>>>(P);
                    else list_fwd_kernel<false, true><<<grid, LIST_THREADS, sm_f, ks>>>(P);
                } else if (kind == 1) {
                    if (b->any_mscc) list_rev_kernel<true><<<grid, LIST_THREADS, sm_2, ks>>>(P);
                    else list_rev_kernel<false><<<grid, LIST_THREADS, 0, ks>>>(P);
                } else {
                    list_edge_kernel<<<grid, LIST_THREADS, sm_2, ks>>>(P);
                }
                launches++;
            }
            if (split && kind < 2 && pass + 1 == n_pass) CU_TRY(c, cudaEventRecord(c->evk[kind], st));   // (per-kernel split: back-to-back passes only)
        }
    }
    if (concurrent) {
        CU_TRY(c, cudaEventRecord(c->ev_side, c->side_stream));
        CU_TRY(c, cudaStreamWaitEvent(st, c->ev_side, 0));
    }
    P.job_lo = 0; P.job_hi = (uint32_t)n;
    list_finalize_kernel<<<dim3((unsigned)n, 4), 256, 0, st>>>(P);
    launches++;
    CU_TRY(c, cudaGetLastError());
    CU_TRY(c, cudaEventRecord(c->ev[2], st));
    b->h2d_pending = false;
    c->timing.launches = launches;
    c->timing.window_launches = launches;
    c->timing.window_word_shift_pairs = 0;
    c->timing.path = 3;
    c->timing_split = split;
    c->timing.build_ms = c->timing.window_ms = c->timing.total_ms = -1.f;
    c->timing.ev_fwd_ms = c->timing.ev_rev_ms = c->timing.ev_edge_ms = 0.f;
    return PMB_OK;
}

// The list path found its inputs unsorted (or the intervals overlapping): the batch moves to
// the event path for good -- bit arrays are planned now, from the inputs already in HBM -- and
// is executed again.
static int execute_events(pmb_ctx* c, pmb_batch* b);
static int fallback_to_events(pmb_ctx* c, pmb_batch* b)
{
    b->path = 2;
    if (!b->arrays_planned) {
        int rc = layout_arrays(c, b);
        if (rc == PMB_OK) rc = plan_arrays(c, b);
        if (rc != PMB_OK) return rc;
    }
    return execute_events(c, b);
}

// ------------------------------------------------------------------ batch: execute (event path)
static int execute_events(pmb_ctx* c, pmb_batch* b)
{
    cudaStream_t st = c->stream;
    uint32_t launches = 0;
    const size_t n = b->jobs.size();
    // PMB_STAGE_TIMING=1: per-stage device times of the build on stderr (debug aid, adds a sync)
    static const bool stage_timing = getenv("PMB_STAGE_TIMING") != nullptr;
    cudaEvent_t sev[8] = {};
    int n_sev = 0;
    auto mark = [&]() {
        if (!stage_timing || n_sev >= 8) return;
        cudaEventCreate(&sev[n_sev]);
        cudaEventRecord(sev[n_sev++], st);
    };

    CU_TRY(c, cudaEventRecord(c->ev[0], st));
    mark();
    // only the scattered arrays need zeroing (F, R of all jobs come first; M too when it is not
    // gathered); M, Er and Or are written in full by mask_build_multi_kernel.  One launch also
    // clears the counters, the edge sums and the flags.
    {
        ZeroSegs z{};
        // F, R unless they are packed; M unless it is gathered (layout: all F, R | all M | Er, Or)
        const size_t z_lo = b->pack_reads ? b->m_begin : 0, z_hi = b->zero_words;
        z.p[0] = (uint4*)(b->d_bits + z_lo);  z.n16[0] = z_hi > z_lo ? (z_hi - z_lo) / 4 : 0;
        z.p[1] = (uint4*)b->d_counters;  z.n16[1] = b->counter_words * 8 / 16;
        z.p[2] = (uint4*)b->d_aux;       z.n16[2] = b->aux_words * 8 / 16;
        z.p[3] = (uint4*)b->d_err;       z.n16[3] = 4 * (4 + 128) / 16;
        zero_segments_kernel<<<c->num_sms * 8, 512, 0, st>>>(z);
        launches++;
    }
    mark();

    auto blocks_for = [&](uint64_t items, int per_block) {
        uint64_t g = (items + per_block - 1) / per_block;
        return (unsigned)std::min<uint64_t>(std::max<uint64_t>(g, 1), (uint64_t)c->num_sms * 16);
    };
    // One pass = build + event kernels over a run of jobs: the whole batch, or (first execute
    // after an asynchronous prepare) one H2D group at a time, each behind its copy event, so the
    // kernels of group g run while the inputs of group g+1 are still crossing PCIe.
    pmb_batch::Group whole;
    whole.job_lo = 0; whole.job_hi = (uint32_t)n;
    whole.sc_lo = 0; whole.sc_hi = b->n_scatter; whole.sc_items = b->scatter_items;
    whole.fl_lo = 0; whole.fl_hi = b->n_fill; whole.fl_items = b->fill_items;
    whole.dv_lo = 0; whole.dv_hi = b->n_derive; whole.dv_ctas = b->derive_ctas;
    whole.tile_lo = 0; whole.tile_hi = b->ev_total_tiles;
    whole.pk_lo = 0; whole.pk_hi = b->n_pack; whole.pk_chunks = b->pack_chunks;
    whole.ed_ctas = b->edge_ctas;
    whole.fr_lo = 0; whole.fr_hi = b->m_begin;
    const bool grouped = b->h2d_pending && b->groups.size() > 1;
    const size_t n_pass = grouped ? b->groups.size() : 1;
    if (b->h2d_pending && !grouped)
        for (auto& gr : b->groups) CU_TRY(c, cudaStreamWaitEvent(st, gr.done, 0));
    if (grouped) {      // the side stream starts behind the memsets above
        CU_TRY(c, cudaEventRecord(c->ev_side, st));
        CU_TRY(c, cudaStreamWaitEvent(c->side_stream, c->ev_side, 0));
    }

    EventParams P{};
    P.jobs = b->d_evjobs;
    P.n_jobs = (uint32_t)n;
    P.tile_words = b->ev_tile_words;
    P.max_shift = b->max_shift;
    P.read_len = b->read_len;
    P.stride = b->stride;
    P.mem_lo = b->d_bits_raw;
    P.mem_hi = b->d_bits + std::max<size_t>(b->bits_words, 4) + EV_PAD_WORDS;
    P.chk = b->d_err;
    if (c->check_selftest) P.mem_hi = P.mem_lo;
    const uint32_t n_cols_total = (b->n_shifts + 31) / 32;
    const uint32_t bps = c->event_blocks_per_sm > 0 ? (uint32_t)c->event_blocks_per_sm : 8;
    uint32_t window_launches = 0;
#define PMB_EV_LAUNCH(KIND, NCC, MSCC)                                                              \
    do {                                                                                            \
        auto k = event_window_kernel<KIND, NCC, MSCC>;                                              \
        const size_t sm = event_smem_bytes<KIND>();                                                 \
        for (uint32_t col0 = 0; col0 < n_cols_total; col0 += 32) {                                  \
            P.col0 = col0;                                                                          \
            P.n_cols = std::min(32u, n_cols_total - col0);                                          \
            k<<<grid, 32 * EV_WARPS, sm, ps>>>(P);                                                  \
            window_launches++;                                                                      \
        }                                                                                           \
    } while (0)
    if (!c->event_attr_done) {      // once per context (the attribute is per device), not before every launch
#define PMB_EV_ATTR(KIND, NCC, MSCC)                                                                                    \
    CU_TRY(c, cudaFuncSetAttribute(event_window_kernel<KIND, NCC, MSCC>, cudaFuncAttributeMaxDynamicSharedMemorySize,   \
                                   (int)event_smem_bytes<KIND>()))
        PMB_EV_ATTR(EV_FWD, true, true); PMB_EV_ATTR(EV_FWD, true, false); PMB_EV_ATTR(EV_FWD, false, true);
        PMB_EV_ATTR(EV_REV, false, true); PMB_EV_ATTR(EV_EDGE, false, true);
#undef PMB_EV_ATTR
        c->event_attr_done = true;
    }
    for (size_t pass = 0; pass < n_pass; pass++) {
        const pmb_batch::Group& G = grouped ? b->groups[pass] : whole;
        // odd groups run on the side stream: a group's kernel tails then overlap the next
        // group's build instead of leaving the SMs idle between 7 short launches
        const bool side = grouped && (pass & 1);
        cudaStream_t ps = side ? c->side_stream : st;
        uint32_t* long_count = b->d_long + (side ? 1 : 0);
        unsigned long long* long_list = b->d_long64 + (side ? b->fill_items + 1 : 0);
        if (grouped) CU_TRY(c, cudaStreamWaitEvent(ps, G.done, 0));
        // F, R (strand totals are counted by the forward-read event kernel).  Packed build: one
        // pass writes every word once; the zero + atomicOr scatter stays behind it, gated on the
        // "reads unsorted" flag of the group.
        int* rgate = b->pack_reads ? b->d_err + 4 + 64 + (grouped ? (int)pass : 0) : nullptr;
        if (b->pack_reads && G.pk_chunks) {
            uint32_t* run_count = b->d_long + (side ? 3 : 2);
            ZeroRun* runs = b->d_runs + (side ? b->runs_cap : 0);
            CU_TRY(c, cudaMemsetAsync(run_count, 0, 4, ps));
            pack_reads_multi_kernel<<<(unsigned)((G.pk_chunks + PACK_WARPS - 1) / PACK_WARPS), 32 * PACK_WARPS, 0, ps>>>(
                b->d_pack + G.pk_lo, b->d_pack_prefix + G.pk_lo, G.pk_hi - G.pk_lo, b->d_err, rgate, runs, run_count,
                (uint32_t)b->runs_cap);
            zero_runs_kernel<<<c->num_sms * 4, 256, 0, ps>>>(b->d_pack + G.pk_lo, runs, run_count, (uint32_t)b->runs_cap);
            launches += 2;
            if (G.sc_items && G.fr_hi > G.fr_lo) {
                zero_gated_kernel<<<c->num_sms * 8, 512, 0, ps>>>((uint4*)(b->d_bits + G.fr_lo), (G.fr_hi - G.fr_lo) / 4, rgate);
                launches++;
            }
        }
        if (G.sc_items) {
            scatter_multi_kernel<<<blocks_for(G.sc_items, 256), 256, 0, ps>>>(
                b->d_scatter + G.sc_lo, b->d_scatter_prefix + G.sc_lo, G.sc_hi - G.sc_lo, b->d_err, rgate);
            launches++;
        }
        if (!grouped) mark();
        // M (+ Er, Or, direct lags).  Gather build: verify that the intervals are sorted and
        // disjoint, then ONE pass writes everything; the scatter-style kernels stay in the stream
        // behind it, gated on the "unsorted" flag (they return at once in the normal case).
        int* gate = b->gather_m ? b->d_err + 4 + (grouped ? (int)pass : 0) : nullptr;   // per group: groups run concurrently
        if (b->gather_m) {
            if (G.fl_items) {
                check_intervals_multi_kernel<<<blocks_for(G.fl_items, 256), 256, 0, ps>>>(
                    b->d_fill + G.fl_lo, b->d_fill_prefix + G.fl_lo, G.fl_hi - G.fl_lo, b->max_shift ? 2 : 1,
                    b->d_err, gate);
                launches++;
            }
            if (G.dv_ctas) {
                mask_build_multi_kernel<<<(unsigned)G.dv_ctas, 256, 0, ps>>>(
                    b->d_derive + G.dv_lo, b->d_derive_prefix + G.dv_lo, G.dv_hi - G.dv_lo, gate);
                launches++;
            }
            if (G.dv_ctas && G.fl_items) {      // fallback: zero this group's M arrays first
                const size_t m_lo = b->jobs_m_lo(G.job_lo, G.job_hi), m_hi = b->jobs_m_hi(G.job_lo, G.job_hi);
                if (m_hi > m_lo) {
                    zero_gated_kernel<<<c->num_sms * 8, 512, 0, ps>>>((uint4*)(b->d_bits + m_lo), (m_hi - m_lo) / 4, gate);
                    launches++;
                }
            }
        }
        if (G.fl_items) {
            CU_TRY(c, cudaMemsetAsync(long_count, 0, 4, ps));
            fill_multi_kernel<<<blocks_for(G.fl_items, 8), 256, 0, ps>>>(
                b->d_fill + G.fl_lo, b->d_fill_prefix + G.fl_lo, G.fl_hi - G.fl_lo, b->d_err, long_list,
                long_count, gate);
            fill_long_multi_kernel<<<c->num_sms * 4, 256, 0, ps>>>(b->d_fill + G.fl_lo, long_list, long_count, gate);
            launches += 2;
        }
        if (!grouped) mark();
        if (G.dv_ctas && (!b->gather_m || G.fl_items)) {
            derive_multi_kernel<<<(unsigned)G.dv_ctas, 256, 0, ps>>>(
                b->d_derive + G.dv_lo, b->d_derive_prefix + G.dv_lo, G.dv_hi - G.dv_lo, b->max_shift ? 2 : 1, gate);
            launches++;
        }
        CU_TRY(c, cudaGetLastError());
        if (!grouped) CU_TRY(c, cudaEventRecord(c->ev[1], st));
        if (!grouped) mark();

        P.tile_lo = G.tile_lo;
        P.tile_hi = G.tile_hi;
        if (G.tile_hi == G.tile_lo) continue;
        const uint32_t grid = std::max(1u, std::min(G.tile_hi - G.tile_lo, (uint32_t)c->num_sms * bps));
        if (b->any_ncc && b->any_mscc) PMB_EV_LAUNCH(EV_FWD, true, true);
        else if (b->any_ncc) PMB_EV_LAUNCH(EV_FWD, true, false);
        else if (b->any_mscc) PMB_EV_LAUNCH(EV_FWD, false, true);
        if (!grouped) CU_TRY(c, cudaEventRecord(c->evk[0], st));
        if (b->any_mscc) PMB_EV_LAUNCH(EV_REV, false, true);
        if (!grouped) CU_TRY(c, cudaEventRecord(c->evk[1], st));
        if (b->any_mscc || b->any_autocorr) {
            // run edges: from the interval list when M was gathered from it; the scan over M stays
            // behind it, gated on the same "unsorted" flag
            const bool listed = b->gather_m && b->edge_list_ok && !getenv("PMB_NO_EDGE_LIST");
            if (listed && G.ed_ctas) {
                for (uint32_t col0 = 0; col0 < n_cols_total; col0 += 32) {
                    EdgeListParams E{col0, std::min(32u, n_cols_total - col0), b->max_shift, b->stride};
                    edge_list_kernel<<<(unsigned)G.ed_ctas, 256, 0, ps>>>(
                        b->d_derive + G.dv_lo, b->d_edge_prefix + G.dv_lo, G.dv_hi - G.dv_lo, E, gate);
                    window_launches++;
                }
            }
            P.gate = listed ? gate : nullptr;
            if (!listed || G.fl_items) PMB_EV_LAUNCH(EV_EDGE, false, true);
            P.gate = nullptr;
        }
    }
#undef PMB_EV_LAUNCH
    if (grouped) {      // join the side stream; the per-kernel split is not meaningful here
        CU_TRY(c, cudaEventRecord(c->ev_side, c->side_stream));
        CU_TRY(c, cudaStreamWaitEvent(st, c->ev_side, 0));
        CU_TRY(c, cudaEventRecord(c->ev[1], st));
        CU_TRY(c, cudaEventRecord(c->evk[0], st));
        CU_TRY(c, cudaEventRecord(c->evk[1], st));
    }
    b->h2d_pending = false;
    if (n_sev > 1) {
        cudaEventSynchronize(sev[n_sev - 1]);
        static const char* names[] = {"memset", "scatter", "mask/fill", "derive", "?", "?", "?"};
        fprintf(stderr, "[pmb stages]");
        for (int i = 1; i < n_sev; i++) {
            float ms = 0; cudaEventElapsedTime(&ms, sev[i - 1], sev[i]);
            fprintf(stderr, " %s %.1f us", names[i - 1], ms * 1e3f);
        }
        fprintf(stderr, "\n");
    }
    for (int i = 0; i < n_sev; i++) cudaEventDestroy(sev[i]);
    if (const char* dump = getenv("PMB_DUMP_ARRAYS")) {      // debug aid: job 0's F, R, M, Er|Or to a file
        const JobPlan& j = b->jobs[0];
        cudaStreamSynchronize(st);
        std::vector<uint32_t> h(4 * j.array_words, 0u);
        if (!j.autocorr) cudaMemcpy(h.data(), b->d_bits + j.bits, 2 * j.array_words * 4, cudaMemcpyDeviceToHost);
        if (j.autocorr || (j.mode & MODE_MSCC))
            cudaMemcpy(h.data() + 2 * j.array_words, b->d_bits + j.bits_m, j.array_words * 4, cudaMemcpyDeviceToHost);
        if (!j.autocorr && (j.mode & MODE_MSCC))
            cudaMemcpy(h.data() + 3 * j.array_words, b->d_bits + j.bits_d, j.array_words * 4, cudaMemcpyDeviceToHost);
        if (FILE* f = fopen(dump, "wb")) { fwrite(h.data(), 4, h.size(), f); fclose(f); }
    }
    if (b->any_mscc || b->any_autocorr) {
        finalize_dmp_kernel<<<(unsigned)n, 256, 0, st>>>(b->d_evjobs, (uint32_t)n, b->stride, b->max_shift,
                                                         b->gather_m ? b->d_err + 4 : nullptr, grouped ? 1 : 0);
        window_launches++;
    }
    CU_TRY(c, cudaGetLastError());
    CU_TRY(c, cudaEventRecord(c->ev[2], st));
    c->timing.launches = launches + window_launches;
    c->timing.window_launches = window_launches;
    c->timing.window_word_shift_pairs = 0;
    c->timing.path = 2;
    c->timing.build_ms = c->timing.window_ms = c->timing.total_ms = -1.f;
    return PMB_OK;
}

static int execute_dense(pmb_ctx* c, pmb_batch* b);

extern "C" int pmb_batch_execute(pmb_ctx* c, pmb_batch* b)
{
    if (!c || !b) return PMB_ERR_ARG;
    return guarded(c, [&]() -> int {
        CU_TRY(c, cudaSetDevice(c->device));
        return b->path == 3 ? execute_lists(c, b) : b->path == 2 ? execute_events(c, b) : execute_dense(c, b);
    });
}

static int execute_dense(pmb_ctx* c, pmb_batch* b)
{
    cudaStream_t st = c->stream;
    const uint32_t H = b->halo_words;
    const int64_t NEG = -(1ll << 40), POS = (1ll << 40);
    uint32_t launches = 0;

    CU_TRY(c, cudaEventRecord(c->ev[0], st));
    if (b->h2d_pending) {       // asynchronous prepare: the inputs arrive on the copy stream
        for (auto& gr : b->groups) CU_TRY(c, cudaStreamWaitEvent(st, gr.done, 0));
        b->h2d_pending = false;
    }
    CU_TRY(c, cudaMemsetAsync(b->d_bits, 0, std::max<size_t>(b->bits_words, 4) * 4, st));
    CU_TRY(c, cudaMemsetAsync(b->d_counters, 0, b->counter_words * 8, st));
    CU_TRY(c, cudaMemsetAsync(b->d_err, 0, 4 * (4 + 128), st));
    const size_t n = b->jobs.size();
    unsigned long long* d_totals = b->d_counters + n * PMB_NCOUNTERS * b->stride;

    auto blocks_for = [&](uint64_t items, int per_block) {
        uint64_t g = (items + per_block - 1) / per_block;
        return (unsigned)std::min<uint64_t>(std::max<uint64_t>(g, 1), (uint64_t)c->num_sms * 16);
    };

    for (size_t i = 0; i < n; i++) {
        JobPlan& j = b->jobs[i];
        Geometry g;
        g.origin_bit = j.origin_word * 32;
        g.halo_bits = (int64_t)H * 32;
        g.total_bits = (int64_t)j.array_words * 32;
        uint32_t* base = b->d_bits + j.bits;
        auto arr = [&](int slot) { return base + (size_t)slot * j.array_words; };

        if (j.autocorr) {
            const int n_arr = arrays_of(j);
            for (int pass = 0; pass < n_arr && j.a.n_intervals; pass++) {
                uint32_t* M = arr(pass);
                int64_t clo = pass == 0 ? NEG : j.lo, chi = pass == 0 ? POS : j.hi;
                CU_TRY(c, cudaMemsetAsync(b->d_long, 0, 4, st));
                fill_intervals_kernel<<<blocks_for(j.a.n_intervals, 8), 256, 0, st>>>(
                    M, g, b->d_in + j.in_mb, b->d_in + j.in_me, j.a.n_intervals, 0, clo, chi, j.nbits,
                    b->d_err, b->d_long + 1, b->d_long);
                fill_long_kernel<<<c->num_sms * 4, 256, 0, st>>>(M, g, b->d_in + j.in_mb,
                                                                 b->d_in + j.in_me, 0, clo, chi,
                                                                 b->d_long + 1, b->d_long);
                launches += 2;
            }
            continue;
        }
        if (j.a.n_fwd) {
            scatter_bits_kernel<<<blocks_for(j.a.n_fwd, 256), 256, 0, st>>>(
                arr(A_F), g, b->d_in + j.in_fwd, j.a.n_fwd, j.lo, j.hi, j.nbits, b->d_err);
            launches++;
        }
        if (j.a.n_rev) {
            scatter_bits_kernel<<<blocks_for(j.a.n_rev, 256), 256, 0, st>>>(
                arr(A_R), g, b->d_in + j.in_rev, j.a.n_rev, NEG, POS, j.nbits, b->d_err);
            launches++;
        }
        if ((j.mode & MODE_MSCC) && j.a.n_intervals) {
            for (int pass = 0; pass < 2; pass++) {
                uint32_t* dst = arr(pass == 0 ? A_M : A_MP);
                int64_t down = pass == 0 ? 0 : (int64_t)b->read_len - 1;
                int64_t clo = pass == 0 ? j.lo : NEG, chi = pass == 0 ? j.hi : POS;
                CU_TRY(c, cudaMemsetAsync(b->d_long, 0, 4, st));
                fill_intervals_kernel<<<blocks_for(j.a.n_intervals, 8), 256, 0, st>>>(
                    dst, g, b->d_in + j.in_mb, b->d_in + j.in_me, j.a.n_intervals, down, clo, chi,
                    j.nbits, b->d_err, b->d_long + 1, b->d_long);
                fill_long_kernel<<<c->num_sms * 4, 256, 0, st>>>(dst, g, b->d_in + j.in_mb,
                                                                 b->d_in + j.in_me, down, clo, chi,
                                                                 b->d_long + 1, b->d_long);
                launches += 2;
            }
        }
        const bool mscc = (j.mode & MODE_MSCC) != 0;
        prep_kernel<<<blocks_for(j.body_words, 1024), 256, 0, st>>>(
            arr(A_F) + H, mscc ? arr(A_M) + H : nullptr, mscc ? arr(A_FM) + H : nullptr, arr(A_R) + H,
            j.body_words, g.origin_bit, j.lo, j.hi, d_totals + 2 * i);
        launches++;
    }
    CU_TRY(c, cudaGetLastError());
    CU_TRY(c, cudaEventRecord(c->ev[1], st));

    // shift-window launches planned by pmb_batch_prepare
    uint32_t window_launches = 0;
    uint64_t pairs = 0;
    for (const WinLaunch& L : b->launches) {
        cudaError_t e;
        int variant = c->variant < 0 ? 1 : c->variant;
        dim3 grid(L.blocks), block(L.warps * 32);
        if (L.mode == MODE_NCC) e = launch_window<MODE_NCC>(c, variant, L.P, grid, block, L.smem);
        else if (L.mode == MODE_MSCC) e = launch_window<MODE_MSCC>(c, variant, L.P, grid, block, L.smem);
        else e = launch_window<MODE_BOTH>(c, variant, L.P, grid, block, L.smem);
        CU_TRY(c, e);
        window_launches++;
        pairs += L.pairs;
    }
    CU_TRY(c, cudaEventRecord(c->ev[2], st));
    c->timing.launches = launches + window_launches;
    c->timing.window_launches = window_launches;
    c->timing.window_word_shift_pairs = pairs;
    c->timing.path = 1;
    c->timing.ev_fwd_ms = c->timing.ev_rev_ms = c->timing.ev_edge_ms = 0.f;
    c->timing.build_ms = c->timing.window_ms = c->timing.total_ms = -1.f;   // resolved lazily
    return PMB_OK;
}

// ------------------------------------------------------------------ batch: fetch
extern "C" int pmb_batch_counters(pmb_ctx* c, pmb_batch* b, void** dev_ptr, uint64_t* n_int64,
                                  uint32_t* stride)
{
    if (!c || !b) return PMB_ERR_ARG;
    if (dev_ptr) *dev_ptr = (void*)b->d_counters;
    if (n_int64) *n_int64 = (uint64_t)b->counter_words;
    if (stride) *stride = b->stride;
    return PMB_OK;
}

// The flags of the batch's LAST execute: the list path alternates between two pairs (see ListJob),
// the other paths clear and use the first.
static int* batch_flags(pmb_batch* b)
{
    return b->d_err + (b->path == 3 && b->list_execs ? 2 * ((b->list_execs - 1) & 1u) : 0);
}

// After an execute on the list path: has the device found the inputs unsorted / overlapping?  Then
// the batch moves to the event path and is executed again (synchronises the stream).
static int settle_path(pmb_ctx* c, pmb_batch* b)
{
    if (b->path != 3 || b->list_checked) return PMB_OK;
    int* h_flags = (int*)((char*)c->h_stage + c->h_stage_bytes - 16);
    CU_TRY(c, cudaMemcpyAsync(h_flags, batch_flags(b), 8, cudaMemcpyDeviceToHost, c->stream));
    CU_TRY(c, cudaStreamSynchronize(c->stream));
    if (h_flags[1] && !h_flags[0]) return fallback_to_events(c, b);
    b->list_checked = true;     // the inputs are resident: what held for this execute holds for the next
    return PMB_OK;
}

static int ensure_stage(pmb_ctx* c, size_t bytes)
{
    if (c->h_stage_bytes >= bytes) return PMB_OK;
    if (c->h_stage) cudaFreeHost(c->h_stage);
    c->h_stage = nullptr;
    c->h_stage_bytes = 0;
    CU_TRY(c, cudaHostAlloc(&c->h_stage, bytes, cudaHostAllocDefault));
    c->h_stage_bytes = bytes;
    return PMB_OK;
}

extern "C" int pmb_batch_validate(pmb_ctx* c, pmb_batch* b)
{
    if (!c || !b) return PMB_ERR_ARG;
    return guarded(c, [&]() -> int {
        CU_TRY(c, cudaSetDevice(c->device));
        int rc = ensure_stage(c, b->counter_words * 8 + 32);
        return rc == PMB_OK ? settle_path(c, b) : rc;
    });
}

extern "C" int pmb_batch_fetch(pmb_ctx* c, pmb_batch* b, pmb_chrom_out* o)
{
    if (!c || !b || !o) return PMB_ERR_ARG;
    CU_TRY(c, cudaSetDevice(c->device));
    const size_t bytes = b->counter_words * 8 + 32;
    {
        int rc = ensure_stage(c, bytes);
        if (rc != PMB_OK) return rc;
    }
    int64_t* h = (int64_t*)c->h_stage;
    int* h_err = (int*)((char*)c->h_stage + b->counter_words * 8);
    CU_TRY(c, cudaMemcpyAsync(h, b->d_counters, b->counter_words * 8, cudaMemcpyDeviceToHost, c->stream));
    CU_TRY(c, cudaMemcpyAsync(h_err, batch_flags(b), 8, cudaMemcpyDeviceToHost, c->stream));
    CU_TRY(c, cudaStreamSynchronize(c->stream));
    if (b->path == 3 && !b->list_checked && h_err[1] && !h_err[0]) {
        // unsorted input on the list path: recompute on the event path, then read again
        int rc = guarded(c, [&]() -> int { return fallback_to_events(c, b); });
        if (rc != PMB_OK) return rc;
        CU_TRY(c, cudaMemcpyAsync(h, b->d_counters, b->counter_words * 8, cudaMemcpyDeviceToHost, c->stream));
        CU_TRY(c, cudaMemcpyAsync(h_err, batch_flags(b), 8, cudaMemcpyDeviceToHost, c->stream));
        CU_TRY(c, cudaStreamSynchronize(c->stream));
    } else if (b->path == 3) {
        b->list_checked = true;
    }
    if (*h_err >= 1000) {       // a -DPMB_CHECKED build only (PMB_CHECK in the kernels)
        c->err = "device-side check failed at source line " + std::to_string(*h_err - 1000);
        return PMB_ERR_CUDA;
    }
    if (*h_err) {
        c->err = "a read position or mappability interval lies beyond chrom_len + read_len + "
                 "max_shift + 100 bits";
        return PMB_ERR_RANGE;
    }
    const size_t n = b->jobs.size(), S = b->n_shifts;
    const int64_t* totals = h + n * PMB_NCOUNTERS * b->stride;
    for (size_t i = 0; i < n; i++) {
        const int64_t* base = h + i * PMB_NCOUNTERS * b->stride;
        const JobPlan& j = b->jobs[i];
        o[i].fwd_total = totals[2 * i];
        o[i].rev_total = totals[2 * i + 1];
        auto put = [&](int64_t* dst, int slot) {
            if (dst) memcpy(dst, base + (size_t)slot * b->stride, S * sizeof(int64_t));
        };
        if (j.autocorr) { put(o[i].dmp, C_DMP); continue; }
        if (j.mode & MODE_NCC) put(o[i].ncc, C_NCC);
        if (j.mode & MODE_MSCC) {
            put(o[i].dmp, C_DMP); put(o[i].mfwd, C_MFWD); put(o[i].mrev, C_MREV); put(o[i].mscc, C_MSCC);
        }
    }
    return PMB_OK;
}

// ------------------------------------------------------------------ one-call forms
extern "C" int pmb_run_batch(pmb_ctx* c, const pmb_chrom_args* a, pmb_chrom_out* o, uint32_t n)
{
    pmb_batch* b = nullptr;
    // the caller's arrays stay valid for the whole call, so the input copy may overlap the kernels
    int rc = guarded(c, [&]() -> int { return prepare_impl(c, a, n, &b, true); });
    if (rc != PMB_OK) return rc;
    rc = pmb_batch_execute(c, b);
    if (rc == PMB_OK) rc = pmb_batch_fetch(c, b, o);
    pmb_batch_free(c, b);
    return rc;
}

extern "C" int pmb_run_chrom(pmb_ctx* c, const pmb_chrom_args* a, pmb_chrom_out* o)
{
    return pmb_run_batch(c, a, o, 1);
}

// ------------------------------------------------------------------ the collective (NCCL over NVLink)
// Position shards add up: every rank carries the same job list, so the counter blocks
// have one layout and ONE ncclAllReduce(int64, sum) of the whole block completes the result on
// every rank.  NCCL is bound at run time: a copy already loaded into the process (torch brings
// one) is preferred, else libnccl.so.2 from the loader's path, or $PMB_NCCL_LIB -- the library
// itself links no NCCL, so single-GPU hosts need none.
struct NcclApi {
    void* handle = nullptr;
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*GroupStart)() = nullptr;
    ncclResult_t (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    std::string error;
};

static NcclApi* nccl_api()
{
    static NcclApi api;
    static bool tried = false;
    if (tried) return &api;
    tried = true;
    const char* env = getenv("PMB_NCCL_LIB");
    void* h = nullptr;
    if (env && *env) h = dlopen(env, RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);      // the process's own copy, if it has one
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!h) { api.error = std::string("NCCL not found (libnccl.so.2; set PMB_NCCL_LIB): ") + (dlerror() ? dlerror() : ""); return &api; }
    api.handle = h;
    bool ok = true;
    auto sym = [&](const char* name) { void* p = dlsym(h, name); if (!p) { ok = false; api.error = std::string("NCCL symbol missing: ") + name; } return p; };
    api.GetUniqueId = (decltype(api.GetUniqueId))sym("ncclGetUniqueId");
    api.CommInitRank = (decltype(api.CommInitRank))sym("ncclCommInitRank");
    api.CommInitAll = (decltype(api.CommInitAll))sym("ncclCommInitAll");
    api.CommDestroy = (decltype(api.CommDestroy))sym("ncclCommDestroy");
    api.AllReduce = (decltype(api.AllReduce))sym("ncclAllReduce");
    api.GroupStart = (decltype(api.GroupStart))sym("ncclGroupStart");
    api.GroupEnd = (decltype(api.GroupEnd))sym("ncclGroupEnd");
    api.GetErrorString = (decltype(api.GetErrorString))sym("ncclGetErrorString");
    if (!ok) api.handle = nullptr;
    return &api;
}

#define NCCL_TRY(ctx, api, call)                                                                  \
    do {                                                                                          \
        ncclResult_t r__ = (call);                                                                \
        if (r__ != ncclSuccess) {                                                                 \
            (ctx)->err = std::string(#call) + " failed: " + (api)->GetErrorString(r__);           \
            return PMB_ERR_CUDA;                                                                  \
        }                                                                                         \
    } while (0)

extern "C" int pmb_comm_get_id(void* id_out)
{
    if (!id_out) return PMB_ERR_ARG;
    NcclApi* api = nccl_api();
    if (!api->handle) { g_create_error = api->error; return PMB_ERR_CUDA; }
    ncclUniqueId id;
    if (api->GetUniqueId(&id) != ncclSuccess) { g_create_error = "ncclGetUniqueId failed"; return PMB_ERR_CUDA; }
    memcpy(id_out, &id, sizeof id);
    return PMB_OK;
}

extern "C" int pmb_comm_init(pmb_ctx* c, const void* id_in, int rank, int world)
{
    if (!c) return PMB_ERR_ARG;
    if (!id_in || world < 1 || rank < 0 || rank >= world) { c->err = "pmb_comm_init: bad rank / world / id"; return PMB_ERR_ARG; }
    if (c->comm) { c->err = "pmb_comm_init: the context already has a communicator"; return PMB_ERR_ARG; }
    NcclApi* api = nccl_api();
    if (!api->handle) { c->err = api->error; return PMB_ERR_CUDA; }
    CU_TRY(c, cudaSetDevice(c->device));
    ncclUniqueId id;
    memcpy(&id, id_in, sizeof id);
    NCCL_TRY(c, api, api->CommInitRank(&c->comm, world, id, rank));
    c->comm_rank = rank;
    c->comm_world = world;
    return PMB_OK;
}

extern "C" int pmb_comm_init_all(pmb_ctx** ctxs, int n)
{
    if (!ctxs || n < 1 || !ctxs[0]) return PMB_ERR_ARG;
    pmb_ctx* c0 = ctxs[0];
    NcclApi* api = nccl_api();
    if (!api->handle) { c0->err = api->error; return PMB_ERR_CUDA; }
    std::vector<int> devs(n);
    std::vector<ncclComm_t> comms(n, nullptr);
    for (int i = 0; i < n; i++) {
        if (!ctxs[i] || ctxs[i]->comm) { c0->err = "pmb_comm_init_all: null context or communicator already set"; return PMB_ERR_ARG; }
        devs[i] = ctxs[i]->device;
        for (int k = 0; k < i; k++)
            if (devs[k] == devs[i]) { c0->err = "pmb_comm_init_all: one context per device"; return PMB_ERR_ARG; }
    }
    NCCL_TRY(c0, api, api->CommInitAll(comms.data(), n, devs.data()));
    for (int i = 0; i < n; i++) { ctxs[i]->comm = comms[i]; ctxs[i]->comm_rank = i; ctxs[i]->comm_world = n; }
    return PMB_OK;
}

extern "C" void pmb_comm_destroy(pmb_ctx* c)
{
    if (!c || !c->comm) return;
    NcclApi* api = nccl_api();
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    if (api->handle) api->CommDestroy(c->comm);
    c->comm = nullptr;
    c->comm_world = 1;
    c->comm_rank = 0;
}

// the rank's status (range error, inputs unsorted) as two int64 behind the counters: summed with them
__global__ void status_words_kernel(const int* __restrict__ flags, unsigned long long* __restrict__ tail, int checked)
{
    tail[0] = (flags[0] != 0) ? 1ull : 0ull;
    tail[1] = (!checked && flags[1] != 0) ? 1ull : 0ull;
}

static int allreduce_enqueue(NcclApi* api, pmb_ctx** ctxs, pmb_batch** bs, int n)
{
    for (int i = 0; i < n; i++) {
        pmb_ctx* c = ctxs[i];
        CU_TRY(c, cudaSetDevice(c->device));
        if (bs[i]->path != 3) {     // (the list path's finalize kernel has written the words already)
            status_words_kernel<<<1, 1, 0, c->stream>>>(bs[i]->d_err, bs[i]->d_counters + bs[i]->counter_words, 1);
            CU_TRY(c, cudaGetLastError());
        }
    }
    if (n > 1) NCCL_TRY(ctxs[0], api, api->GroupStart());
    for (int i = 0; i < n; i++) {
        pmb_ctx* c = ctxs[i];
        if (n > 1) cudaSetDevice(c->device);
        NCCL_TRY(c, api, api->AllReduce(bs[i]->d_counters, bs[i]->d_counters, bs[i]->counter_words + 2, ncclInt64, ncclSum,
                                        c->comm, c->stream));
    }
    if (n > 1) NCCL_TRY(ctxs[0], api, api->GroupEnd());
    return PMB_OK;
}

// In-place sum of the counter blocks over the communicator, on each context's stream.  The FIRST
// all-reduce of a batch also agrees on the ranks' status (it waits for the result once): a
// position out of range on any rank fails the call on every rank; inputs that the list path found
// unsorted on any rank make every rank recompute its share (that rank on the event path) and
// reduce again.  Later all-reduces of the batch are asynchronous.
extern "C" int pmb_allreduce(pmb_ctx** ctxs, pmb_batch** bs, int n)
{
    if (!ctxs || !bs || n < 1 || !ctxs[0]) return PMB_ERR_ARG;
    pmb_ctx* c0 = ctxs[0];
    return guarded(c0, [&]() -> int {
        NcclApi* api = nccl_api();
        if (!api->handle) { c0->err = api->error; return PMB_ERR_CUDA; }
        for (int i = 0; i < n; i++) {
            if (!ctxs[i] || !bs[i] || !ctxs[i]->comm) { c0->err = "pmb_allreduce: context without a communicator (pmb_comm_init)"; return PMB_ERR_ARG; }
            if (bs[i]->counter_words != bs[0]->counter_words) { c0->err = "pmb_allreduce: the batches' job lists differ"; return PMB_ERR_ARG; }
        }
        int rc = allreduce_enqueue(api, ctxs, bs, n);
        if (rc != PMB_OK) return rc;
        bool all_checked = true;
        for (int i = 0; i < n; i++) all_checked = all_checked && bs[i]->reduce_checked;
        if (all_checked) return PMB_OK;
        // first reduce: read the summed status words (the same on every rank)
        unsigned long long st[2] = {0, 0};
        for (int i = 0; i < n; i++) {
            pmb_ctx* c = ctxs[i];
            CU_TRY(c, cudaSetDevice(c->device));
            unsigned long long s[2];
            CU_TRY(c, cudaMemcpyAsync(s, bs[i]->d_counters + bs[i]->counter_words, 16, cudaMemcpyDeviceToHost, c->stream));
            CU_TRY(c, cudaStreamSynchronize(c->stream));
            st[0] = s[0]; st[1] = s[1];
        }
        if (st[0]) {
            for (int i = 0; i < n; i++)
                ctxs[i]->err = "a read position or mappability interval lies beyond chrom_len + read_len + max_shift + 100 bits "
                               "(on at least one rank)";
            return PMB_ERR_RANGE;
        }
        if (st[1]) {        // some rank's lists were unsorted: every rank recomputes (the sum overwrote its share)
            for (int i = 0; i < n; i++) {
                pmb_ctx* c = ctxs[i];
                CU_TRY(c, cudaSetDevice(c->device));
                int h_flags[2] = {0, 0};
                int rc2;
                if (bs[i]->path == 3 && !bs[i]->list_checked) {
                    // (the flags were cleared by nothing since the execute: still the rank's own)
                    CU_TRY(c, cudaMemcpy(h_flags, batch_flags(bs[i]), 8, cudaMemcpyDeviceToHost));
                    rc2 = h_flags[1] ? fallback_to_events(c, bs[i]) : pmb_batch_execute(c, bs[i]);
                } else {
                    rc2 = pmb_batch_execute(c, bs[i]);
                }
                if (rc2 != PMB_OK) return rc2;
                bs[i]->list_checked = true;
            }
            for (int i = 0; i < n; i++) bs[i]->reduce_checked = true;
            return allreduce_enqueue(api, ctxs, bs, n);
        }
        for (int i = 0; i < n; i++) { bs[i]->reduce_checked = true; bs[i]->list_checked = true; }
        return PMB_OK;
    });
}

// ------------------------------------------------------------------ one chromosome, no allocation
// pmb_run_chrom_noalloc (include/pymasc_b200.h): the list kernels on one chromosome whose lists
// are already in device memory, with every byte of scratch in the array below.  In HOST code the
// array's name is only a stand-in for the symbol -- not a device address: the address comes from
// cudaGetSymbolAddress, once per context, with the context's device current.  Device code names
// the array directly (noalloc_setup_kernel).
//   words [0, 32)    the job table: one ListJob
//   words [32, 64)   entry prefixes of the three kernels: {0, n_fwd}, {0, n_rev}, {0, n_intervals}
//   words [64, ...)  accumulators (ONE copy, cleared by every call): [6][S] pair and difference histograms |
//                    [2] strand totals | [2][S] run-edge products | [2] lags,  S = max_shift + 1 <= PMB_NOALLOC_MAX_SHIFT + 1
constexpr size_t NOALLOC_TABLE_WORDS = 32, NOALLOC_PREFIX_WORDS = 32;
constexpr size_t NOALLOC_ACC_WORDS = (LA_N + 2) * ((size_t)PMB_NOALLOC_MAX_SHIFT + 1) + 4;
constexpr size_t NOALLOC_SCRATCH_WORDS = NOALLOC_TABLE_WORDS + NOALLOC_PREFIX_WORDS + NOALLOC_ACC_WORDS;
static_assert(sizeof(ListJob) <= NOALLOC_TABLE_WORDS * 8, "job table slot too small");
static __device__ __align__(256) unsigned long long g_noalloc_scratch[NOALLOC_SCRATCH_WORDS];

// Writes the job table and the prefixes, and clears what the list kernels accumulate into: the
// accumulators in the scratch array (they hold the previous call's sums) and the caller's output.
__global__ void __launch_bounds__(512) noalloc_setup_kernel(ListJob J, unsigned long long n_fwd,
                                                            unsigned long long n_rev, unsigned long long n_iv,
                                                            unsigned long long* __restrict__ out, uint32_t n_out,
                                                            uint32_t n_acc)
{
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        *reinterpret_cast<ListJob*>(g_noalloc_scratch) = J;
        unsigned long long* pre = g_noalloc_scratch + NOALLOC_TABLE_WORDS;
        pre[0] = 0; pre[1] = n_fwd; pre[2] = 0; pre[3] = n_rev; pre[4] = 0; pre[5] = n_iv;
    }
    unsigned long long* acc = g_noalloc_scratch + NOALLOC_TABLE_WORDS + NOALLOC_PREFIX_WORDS;
    const uint32_t tid = blockIdx.x * blockDim.x + threadIdx.x, nth = gridDim.x * blockDim.x;
    for (uint32_t i = tid; i < n_acc; i += nth) acc[i] = 0;
    for (uint32_t i = tid; i < n_out; i += nth) out[i] = 0;
}

extern "C" int pmb_run_chrom_noalloc(pmb_ctx* c, const pmb_chrom_args* a, int64_t* d_out, uint64_t out_int64)
{
    if (!c) return PMB_ERR_ARG;
    if (!a || !d_out) { c->err = "pmb_run_chrom_noalloc: null argument"; return PMB_ERR_ARG; }
    if (a->max_shift > PMB_NOALLOC_MAX_SHIFT) {
        c->err = "pmb_run_chrom_noalloc: max_shift " + std::to_string(a->max_shift) + " exceeds PMB_NOALLOC_MAX_SHIFT (" +
                 std::to_string(PMB_NOALLOC_MAX_SHIFT) + "): the static scratch array holds no more; use pmb_run_chrom";
        return PMB_ERR_ARG;
    }
    const uint32_t D = a->max_shift, S = D + 1;
    if (out_int64 < PMB_NOALLOC_OUT_INT64(D)) {
        c->err = "pmb_run_chrom_noalloc: d_out holds " + std::to_string(out_int64) + " int64, " +
                 std::to_string(PMB_NOALLOC_OUT_INT64(D)) + " needed";
        return PMB_ERR_ARG;
    }
    if (((uintptr_t)d_out & 7) || ((uintptr_t)a->fwd_pos & 3) || ((uintptr_t)a->rev_pos & 3) ||
        ((uintptr_t)a->map_begin & 3) || ((uintptr_t)a->map_end & 3)) {
        c->err = "pmb_run_chrom_noalloc: misaligned device pointer"; return PMB_ERR_ARG;
    }
    const bool autocorr = (a->flags & PMB_FLAG_AUTOCORR) != 0;
    const bool ncc = !autocorr && (a->flags & PMB_FLAG_NCC), mscc = !autocorr && (a->flags & PMB_FLAG_MSCC);
    if ((a->flags & PMB_FLAG_SAME_TRACK) || (!autocorr && !ncc && !mscc)) {
        c->err = "pmb_run_chrom_noalloc: flags must be NCC and/or MSCC, or AUTOCORR alone"; return PMB_ERR_ARG;
    }
    if (a->read_len < 2) { c->err = "pmb_run_chrom_noalloc: read_len must be >= 2"; return PMB_ERR_ARG; }
    const uint64_t nbits = a->chrom_len + a->read_len + a->max_shift + PMB_EXTRA_ALLOCATE_SIZE;
    if (nbits + 2ull * (a->read_len + a->max_shift) + 64 >= (1ull << 32) || a->n_fwd >= (1ull << 31) || a->n_rev >= (1ull << 31) ||
        a->n_intervals >= (1ull << 31)) {
        c->err = "pmb_run_chrom_noalloc: chromosome or list too long for 32-bit positions"; return PMB_ERR_ARG;
    }
    if ((!autocorr && ((a->n_fwd && !a->fwd_pos) || (a->n_rev && !a->rev_pos))) ||
        ((mscc || autocorr) && a->n_intervals && (!a->map_begin || !a->map_end))) {
        c->err = "pmb_run_chrom_noalloc: list pointer missing"; return PMB_ERR_ARG;
    }
    CU_TRY(c, cudaSetDevice(c->device));
    if (!c->noalloc_scratch) {      // set-up, once per context: the array's address on THIS device
        unsigned long long* p = nullptr;
        CU_TRY(c, cudaGetSymbolAddress((void**)&p, g_noalloc_scratch));
        if (!p) { c->err = "pmb_run_chrom_noalloc: cudaGetSymbolAddress returned no address"; return PMB_ERR_CUDA; }
        c->noalloc_scratch = p;
    }
    list_attributes(c);
    unsigned long long* scratch = c->noalloc_scratch;
    unsigned long long* acc = scratch + NOALLOC_TABLE_WORDS + NOALLOC_PREFIX_WORDS;
    unsigned long long* out = (unsigned long long*)d_out;

    ListJob J{};
    if (!autocorr) {
        J.fwd = a->fwd_pos; J.n_fwd = (uint32_t)a->n_fwd;
        J.rev = a->rev_pos; J.n_rev = (uint32_t)a->n_rev;
        J.flags = (ncc ? EVF_NCC : 0) | (mscc ? EVF_MSCC : 0);
    } else {
        J.flags = EVF_AUTOCORR;
    }
    if (mscc || autocorr) { J.begins = a->map_begin; J.ends = a->map_end; J.n_iv = (uint32_t)a->n_intervals; }
    J.counters = out;                               // [5][S]
    J.totals = out + (size_t)PMB_NCOUNTERS * S;     // [2]
    J.acc = acc;                                    // [LA_N][S] + [2]
    J.aux = acc + (size_t)LA_N * S + 2;             // [2][S] + [2]
    J.nbits = nbits;
    J.lo = (long long)a->range_lo;
    J.hi = a->range_hi ? (long long)std::min<uint64_t>(a->range_hi, nbits) : (long long)nbits;
    if (J.lo > J.hi) J.lo = J.hi;

    cudaStream_t st = c->stream;
    noalloc_setup_kernel<<<c->num_sms, 512, 0, st>>>(J, J.n_fwd, J.n_rev, J.n_iv, out,
                                                     (uint32_t)PMB_NOALLOC_OUT_INT64(D), (uint32_t)((LA_N + 2) * (size_t)S + 4));
    ListParams P{};
    P.jobs = reinterpret_cast<const ListJob*>(scratch);
    P.job_lo = 0; P.job_hi = 1;
    P.max_shift = D; P.read_len = a->read_len; P.stride = S;
    P.pair_cap = LIST_PAIR_CAP;
    P.flags = (int*)(out + (size_t)PMB_NCOUNTERS * S + 2);      // the status word of d_out: {range, unsorted}
    P.acc_off = P.acc_other = 0;                                // one copy: the set-up kernel above clears it
    P.flags_other = nullptr;
    P.status_tail = nullptr;
    const uint32_t bps = c->list_blocks_per_sm > 0 ? (uint32_t)c->list_blocks_per_sm : (uint32_t)LIST_MIN_BLOCKS;
    auto grid_for = [&](unsigned long long items) {
        return (unsigned)std::max<unsigned long long>(1, std::min<unsigned long long>((items + 2047) / 2048,
                                                                                      (unsigned long long)c->num_sms * bps));
    };
    const size_t sm_f = list_smem_bytes(D, list_fwd_hists(mscc)), sm_2 = list_smem_bytes(D, 2);
    const unsigned long long* pre = scratch + NOALLOC_TABLE_WORDS;
    if (J.n_fwd) {
        P.prefix = pre;
        if (ncc && mscc) list_fwd_kernel<true, true><<<grid_for(J.n_fwd), LIST_THREADS, sm_f, st>>>(P);
        else if (ncc) list_fwd_kernel<true, false><<<grid_for(J.n_fwd), LIST_THREADS, sm_f, st>>>(P);
        else list_fwd_kernel<false, true><<<grid_for(J.n_fwd), LIST_THREADS, sm_f, st>>>(P);
    }
    if (J.n_rev) {
        P.prefix = pre + 2;
        if (mscc) list_rev_kernel<true><<<grid_for(J.n_rev), LIST_THREADS, sm_2, st>>>(P);
        else list_rev_kernel<false><<<grid_for(J.n_rev), LIST_THREADS, 0, st>>>(P);
    }
    if (J.n_iv) {
        P.prefix = pre + 4;
        list_edge_kernel<<<grid_for(J.n_iv), LIST_THREADS, sm_2, st>>>(P);
    }
    list_finalize_kernel<<<dim3(1, 4), 256, 0, st>>>(P);
    CU_TRY(c, cudaGetLastError());
    return PMB_OK;
}

// ------------------------------------------------------------------ pipe calibration
extern "C" int pmb_pipe_probe(pmb_ctx* c, int what, double* ops_per_sec, float* ms_out)
{
    if (!c || !ops_per_sec) return PMB_ERR_ARG;
    CU_TRY(c, cudaSetDevice(c->device));
    uint32_t* d_sink = nullptr;
    CU_TRY(c, cudaMallocAsync((void**)&d_sink, 4096, c->stream));
    const int blocks = c->num_sms * 4, threads = 512, iters = 4096;
    double ops_per_thread = 0;
    float best = 1e30f;
    for (int rep = 0; rep < 4; rep++) {
        CU_TRY(c, cudaEventRecord(c->ev[0], c->stream));
        ops_per_thread = launch_probe(what, blocks, threads, iters, d_sink, c->stream);
        CU_TRY(c, cudaGetLastError());
        CU_TRY(c, cudaEventRecord(c->ev[3], c->stream));
        CU_TRY(c, cudaEventSynchronize(c->ev[3]));
        float ms = 0;
        CU_TRY(c, cudaEventElapsedTime(&ms, c->ev[0], c->ev[3]));
        if (rep > 0) best = std::min(best, ms);
    }
    CU_TRY(c, cudaFreeAsync(d_sink, c->stream));
    if (ops_per_thread <= 0) { c->err = "unknown probe"; return PMB_ERR_ARG; }
    *ops_per_sec = ops_per_thread * (double)blocks * threads / (best * 1e-3);
    if (ms_out) *ms_out = best;
    return PMB_OK;
}
