/*
 * pymasc_b200.h -- C ABI of libpymasc_b200.so, the sm_100a replacement for the
 * native layer under PyMaSC's bit-array calculator.
 *
 * Plain pointers and sizes only; no Python, torch or C++ types cross this boundary.
 * Every function returns PMB_OK (0) or a negative pmb_status; nothing aborts or throws.
 * Input arrays are borrowed for the duration of the call.  A pmb_ctx owns one CUDA
 * device, one stream and a device arena; it is not thread-safe, but different contexts
 * (devices) may be driven from different host threads (ctypes drops the GIL).
 *
 * There is NO CPU fallback behind this ABI: without a usable CUDA device pmb_create
 * fails with PMB_ERR_CUDA.
 *
 * Reference interfaces replaced (paths relative to the PyMaSC checkout):
 *   pmb_run_chrom / pmb_run_batch .... CCBitArrayCalculator._calc_correlation,
 *                                      PyMaSC/core/bitarray/mscc.pyx:217-325, i.e. the
 *                                      bit_array_* calls it makes through
 *                                      PyMaSC/core/bitarray/bitarray.pxd:20-36:
 *        fwd_pos / rev_pos scatter ... bit_array_set_bit        external/BitArray/bit_array.c:668
 *                                      (bitarray.__setitem__,   bitarray.pyx:72-79)
 *        map_begin / map_end fill .... bit_array_set_region     bit_array.c:823 (-> :423-479)
 *                                      (_load_mappability,      mscc.pyx:327-349)
 *        fwd_total / rev_total ....... bit_array_num_bits_set   bit_array.c:962-978
 *        ncc[k] ...................... bitarray.acount + rshift bitarray.pyx:109-133, bit_array.c:1993-2011
 *        dmp/mfwd/mrev/mscc[k] ....... bit_array_and x3, count x3, acount, lshift
 *                                      mscc.pyx:291-310, bit_array.c:1614-1636, :1953-1970
 *   PMB_FLAG_AUTOCORR ................ _fill_result's mappable-length loop, mscc.pyx:207-215
 *   popcount ......................... popcnt macro, external/bitcnt.h:23-44
 */
#ifndef PYMASC_B200_H
#define PYMASC_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct pmb_ctx pmb_ctx;
typedef struct pmb_batch pmb_batch;

typedef enum {
    PMB_OK = 0,
    PMB_ERR_CUDA = -1,      /* CUDA runtime error; text in pmb_last_error            */
    PMB_ERR_RANGE = -2,     /* a position / interval lies beyond the bit array       */
                            /* (the reference aborts on a BitArray assert here)      */
    PMB_ERR_ARG = -3,       /* invalid argument                                      */
    PMB_ERR_NOMEM = -4      /* device or pinned-host allocation failed               */
} pmb_status;

/* flags of pmb_chrom_args */
#define PMB_FLAG_NCC      1u  /* compute ncc[k] and the strand totals (not skip_ncc)   */
#define PMB_FLAG_MSCC     2u  /* compute dmp/mfwd/mrev/mscc[k]; needs intervals        */
#define PMB_FLAG_AUTOCORR 4u  /* read-less chromosome: dmp[k] = sum_j M[j] M[j+k]      */
                              /* (mscc.pyx:207-215); other outputs untouched           */
#define PMB_FLAG_SAME_TRACK 8u /* this MSCC job reads the mappability track of the PREVIOUS */
                              /* job of the batch (same chrom_len and range, another read  */
                              /* set: a multi-file run over one BigWig).  map_begin/map_end */
                              /* are ignored; M and dmp[k] (popc(D_k), independent of the  */
                              /* reads) are built once per run of such jobs and every job  */
                              /* of the run reports the same dmp.                          */

#define PMB_EXTRA_ALLOCATE_SIZE 100 /* mscc.pyx:118 */

/*
 * One chromosome (or one position shard of it).
 * Bit array length, as in the reference (mscc.pyx:134,:165):
 *     nbits = chrom_len + read_len + max_shift + 100
 * Positions are 1-based bit indices: fwd_pos = leftmost base of a forward read,
 * rev_pos = pos + query_length - 1 of a reverse read (mscc.pyx:416-417).  Duplicates
 * are allowed and collapse (a bit is set once).  Order is free.
 * Intervals are BigWig-style 0-based half-open (begin, end) with value >= 1.0 already
 * selected by the host; bits begin+1 .. end are set (mscc.pyx:343-344).
 * Counts are taken over j in [range_lo, range_hi); range_hi == 0 means the whole
 * array [0, nbits).  A shard still receives every read/interval (or at least those
 * within max_shift + read_len of its range): halos are built on the device.
 * All chromosomes of one batch must share read_len and max_shift.
 */
typedef struct {
    uint64_t chrom_len;
    uint32_t read_len;
    uint32_t max_shift;
    uint32_t flags;
    uint32_t reserved;
    const uint32_t *fwd_pos;   uint64_t n_fwd;
    const uint32_t *rev_pos;   uint64_t n_rev;
    const uint32_t *map_begin; const uint32_t *map_end; uint64_t n_intervals;
    uint64_t range_lo, range_hi;
} pmb_chrom_args;

/* Caller-allocated outputs; each vector holds max_shift+1 int64.  Vectors that the
 * flags do not request may be NULL. */
typedef struct {
    int64_t fwd_total;   /* popcount(F) over the range   (mscc.pyx:236) */
    int64_t rev_total;   /* popcount(R) over the range   (mscc.pyx:237) */
    int64_t *ncc;        /* sum_j F[j] R[j+k]                     (:314) */
    int64_t *mscc;       /* sum_j F[j] R[j+k] D_k[j]              (:305) */
    int64_t *mfwd;       /* sum_j F[j] D_k[j]                     (:303) */
    int64_t *mrev;       /* sum_j R[j+k] D_k[j]                   (:304) */
    int64_t *dmp;        /* sum_j D_k[j], D_k[j] = M[j] M[j+read_len-1-k]  (:291-298);
                            with PMB_FLAG_AUTOCORR: sum_j M[j] M[j+k]            */
} pmb_chrom_out;

/* Device-side timings of the last pmb_batch_execute, CUDA events on the ctx stream. */
typedef struct {
    float build_ms;     /* memset + scatter + interval fill + prep           */
    float window_ms;    /* shift-window kernel(s)                            */
    float total_ms;     /* first launch to last launch of the execute        */
    uint32_t launches;  /* kernels launched by the execute                   */
    uint32_t window_launches;
    uint64_t window_word_shift_pairs; /* 32-bit word x shift pairs evaluated (dense path) */
    uint32_t path;      /* algorithm the execute used: 1 dense shift-window, 2 event-window, 3 lists */
    float ev_fwd_ms;    /* event / list path: forward-read kernel(s)   (list path: with option "kernel_timing" 1) */
    float ev_rev_ms;    /* event / list path: reverse-read kernel(s)          */
    float ev_edge_ms;   /* event / list path: run-edge kernel(s) + finalize   */
} pmb_timing;

int  pmb_device_count(void);                 /* usable CUDA devices (0 on any error)   */
int  pmb_create(pmb_ctx **out, int device_ordinal);
void pmb_destroy(pmb_ctx *ctx);
const char *pmb_last_error(pmb_ctx *ctx);   /* ctx may be NULL: last create error */

/* Whole path for one chromosome / a list of chromosomes from HOST arrays:
 * H2D, bit-array build, correlation kernels, D2H.  Synchronous for the caller; inside,
 * the input copy runs on a copy stream in "h2d_groups" groups of chromosomes and the
 * kernels of one group overlap the copy of the next. */
int pmb_run_chrom(pmb_ctx *ctx, const pmb_chrom_args *a, pmb_chrom_out *o);
int pmb_run_batch(pmb_ctx *ctx, const pmb_chrom_args *a, pmb_chrom_out *o, uint32_t n);

/* The same path in three steps, so that inputs can stay resident in HBM and the
 * counters can be reduced across GPUs before they leave the device:
 *   prepare : copy positions/intervals to the device (H2D), plan tiles.  With option
 *             "async_h2d" = 1 it returns while the copy is still in flight (grouped as in
 *             pmb_run_batch); the host arrays must then stay valid and unchanged until
 *             the FIRST execute of the batch has completed (pmb_batch_fetch or
 *             pmb_synchronize after it).  Default 0: the arrays are free on return.
 *   execute : build bit arrays + shift-window kernel; counters stay on the device;
 *             may be called repeatedly (each call recomputes from the resident inputs)
 *   fetch   : D2H of the counters into caller-allocated outputs
 * pmb_batch_counters exposes the device counter block for an NCCL all-reduce:
 * int64[n_chroms][PMB_NCOUNTERS][stride], counter order ncc, dmp, mfwd, mrev, mscc,
 * followed by int64[n_chroms][2] strand totals (fwd, rev). */
#define PMB_NCOUNTERS 5
int  pmb_batch_prepare(pmb_ctx *ctx, const pmb_chrom_args *a, uint32_t n, pmb_batch **out);
int  pmb_batch_execute(pmb_ctx *ctx, pmb_batch *b);
int  pmb_batch_fetch(pmb_ctx *ctx, pmb_batch *b, pmb_chrom_out *o);
int  pmb_batch_counters(pmb_ctx *ctx, pmb_batch *b, void **dev_ptr, uint64_t *n_int64,
                        uint32_t *stride);
/* The list path (see "path" below) takes sorted reads and sorted, disjoint intervals -- what the
 * reference's feed rules guarantee (ReadUnsortedError, mscc.pyx:351-366) and what a BigWig holds
 * -- and verifies it on the device while it counts.  pmb_batch_fetch handles the other case by
 * itself (the batch is recomputed on the event path).  A caller that reads the device counters
 * WITHOUT a fetch (an all-reduce on pmb_batch_counters) calls pmb_batch_validate once after the
 * batch's first execute: it waits for the stream and, if the inputs turned out unsorted, moves
 * the batch to the event path and executes it again.  Later executes of the batch need no check. */
int  pmb_batch_validate(pmb_ctx *ctx, pmb_batch *b);
void pmb_batch_free(pmb_ctx *ctx, pmb_batch *b);
int  pmb_get_timing(pmb_ctx *ctx, pmb_timing *t);

/* Stream of the context (cudaStream_t as void*), for callers that order their own
 * work (e.g. an NCCL all-reduce issued through torch) after pmb_batch_execute. */
void *pmb_stream(pmb_ctx *ctx);
int   pmb_synchronize(pmb_ctx *ctx);
/* Run on a stream the caller owns (cudaStream_t as void*; NULL = back to the context's
 * own stream).  The caller keeps the stream alive while the context uses it. */
int   pmb_set_stream(pmb_ctx *ctx, void *stream);

/* Pinned host memory for callers without another way to get it. */
int  pmb_host_alloc(void **out, uint64_t bytes);
void pmb_host_free(void *p);

/* Options (0 = default).  "path": 0 auto, 1 dense shift-window kernel, 2 event-window
 * kernels, 3 list kernels -- three algorithms for the same counts (identical results):
 *   dense  every (bit position, shift) pair from bit arrays in shared memory; cost ~ bits x shifts
 *   events one max_shift-bit window of a bit array per read / run edge;   cost ~ events x shifts/32
 *   lists  no bit arrays: pairs of sorted list entries within max_shift;  cost ~ entries + pairs
 * auto takes the list path while a forward read expects fewer than "list_pairs_limit"
 * (default 64) reverse reads per 1024 shifts, else the event path unless the inputs hold more
 * than 4 events (reads + run edges) per 32-bit word.  Tuning: "list_blocks_per_sm", "list_concurrent"
 * (-1 auto / 0 / 1: the reverse and edge list kernels beside the forward kernel on a second stream; auto does
 * so below 8 M list entries, where a launch is mostly ramp and tail), "list_edge_items" (intervals per CTA of the
 * list path's run-edge kernel at least, default 512), "list_pair_cap" (forward reads between two flushes of the
 * forward list kernel's packed ncc | mscc bins, default and maximum 65535: tests shrink it), "kernel_timing" (list path, 0/1, default 0: with 1 an execute
 * records events between its kernels, so that pmb_get_timing splits the time by kernel -- ev_fwd_ms, ev_rev_ms,
 * ev_edge_ms, zero otherwise -- at ~8 us per step: a kernel then cannot start under the tail of the one before),
 * "variant" (carry-save depth of the shift-window counters), "tile_words",
 * "grid_blocks", "block_warps", "event_tile_words", "event_blocks_per_sm".
 * Input copy: "async_h2d" (0/1, see pmb_batch_prepare), "h2d_groups" (1..64, default 4: at most that many),
 * "h2d_min_group_bytes" (default 16 MiB: no group smaller than that on average; 0 = always h2d_groups),
 * "h2d_growth_x10" (size ratio of consecutive groups in tenths, default 13; the list path uses equal groups, up to
 * twice "h2d_groups" of them at half the minimum size: its kernels outrun the copy).
 * Debug: "check_selftest" (a -DPMB_CHECKED build only: makes every bounds check of the event kernels fail). */
int  pmb_set_option(pmb_ctx *ctx, const char *name, int64_t value);

/*
 * The collective for position shards (the reference has no counterpart: its
 * `-p N` workers exchange pickled results through a multiprocessing.Queue,
 * PyMaSC/handler/calc.py:163-235).  Every rank runs the SAME job list with its own range_lo /
 * range_hi (see pymasc_b200/sharding.py), so the counter blocks share one layout and one
 * ncclAllReduce(int64, sum) over NVLink completes every per-shift vector on every rank.
 * NCCL is bound at run time (a copy already in the process, else libnccl.so.2, or $PMB_NCCL_LIB).
 *
 *   one process per GPU (torchrun, MPI):  rank 0 calls pmb_comm_get_id and hands the
 *       PMB_COMM_ID_BYTES to the other ranks by any side channel; every rank calls pmb_comm_init.
 *   one process, n GPUs:  pmb_comm_init_all over n contexts on n different devices.
 *   pmb_allreduce(ctxs, batches, n): after pmb_batch_execute on every rank, in-place sum of the
 *       counter blocks (what pmb_batch_counters exposes, plus two status words) on each context's
 *       stream; n = 1 in a process per GPU, n = all contexts in one process.  pmb_batch_fetch then
 *       returns the complete vectors on every rank.  The FIRST all-reduce of a batch also agrees
 *       on the ranks' status and waits for the result once: a position beyond the array on any
 *       rank -> PMB_ERR_RANGE on every rank; lists that one rank found unsorted -> every rank
 *       recomputes its share and the reduce is repeated.  Later all-reduces of the same batch
 *       (repeated executes) are asynchronous.
 *   A rank that fails BEFORE the collective (pmb_batch_prepare / execute returned an error) must
 *   not simply leave: the others would wait in NCCL.  Agree on the local status first (the
 *   Python engines do, pymasc_b200/distributed.py).
 */
#define PMB_COMM_ID_BYTES 128
int  pmb_comm_get_id(void *id_out);                                  /* ncclGetUniqueId                 */
int  pmb_comm_init(pmb_ctx *ctx, const void *id, int rank, int world);
int  pmb_comm_init_all(pmb_ctx **ctxs, int n);
int  pmb_allreduce(pmb_ctx **ctxs, pmb_batch **batches, int n);
void pmb_comm_destroy(pmb_ctx *ctx);                                 /* also done by pmb_destroy        */

/*
 * pmb_run_chrom_noalloc -- one chromosome through the list kernels WITHOUT any allocation.
 * (No counterpart in the reference: added for callers
 * that cannot call cudaMalloc -- code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it.)
 *
 * The list path's scratch (job table, difference histograms of mfwd / mrev, run-edge products,
 * the two direct lags: ~257 KB) lives in a file-scope `static __device__` array of the library,
 * sized for max_shift <= PMB_NOALLOC_MAX_SHIFT; every GPU that loads the library holds it until
 * the process exits.  The call issues kernels on the context's stream and returns without
 * waiting; it calls no cudaMalloc / cudaFree / cudaHostAlloc and copies nothing.
 *
 *   a      as for pmb_run_chrom, but fwd_pos / rev_pos / map_begin / map_end are DEVICE pointers
 *          (4-byte aligned), and the lists must be what the reference's feed rules produce:
 *          reads sorted, intervals sorted and disjoint (duplicates allowed).  Flags: PMB_FLAG_NCC,
 *          PMB_FLAG_MSCC, or PMB_FLAG_AUTOCORR alone; read_len >= 2; PMB_FLAG_SAME_TRACK is refused.
 *   d_out  DEVICE pointer, 8-byte aligned, to PMB_NOALLOC_OUT_INT64(max_shift) int64:
 *            [PMB_NCOUNTERS][max_shift+1]  ncc, dmp, mfwd, mrev, mscc (vectors the flags do not
 *                                          select are zero)
 *            [2]                           fwd_total, rev_total
 *            [1]                           status, two int32: low word != 0 -> a position or
 *                                          interval lies beyond the bit array (PMB_ERR_RANGE of
 *                                          the other entry points); high word != 0 -> the lists
 *                                          are not sorted / disjoint and the counts are invalid
 *                                          (use pmb_run_chrom, which handles such input)
 *   out_int64  capacity of d_out in int64; less than PMB_NOALLOC_OUT_INT64(max_shift) is refused
 * Returns PMB_ERR_ARG (text in pmb_last_error) for max_shift > PMB_NOALLOC_MAX_SHIFT and the
 * other refusals above.  The first call on a context looks up the scratch array's device address
 * (cudaGetSymbolAddress) and raises the kernels' shared-memory limit: make that call outside
 * graph capture.
 * ALL calls on one GPU share the scratch array, and it keeps the last call's contents: every call
 * clears what it accumulates into (scratch and d_out) first, and calls that overlap on the GPU --
 * two contexts / streams on one device -- race.  Calls on one context's stream are ordered.
 */
#define PMB_NOALLOC_MAX_SHIFT 4095
#define PMB_NOALLOC_OUT_INT64(max_shift) (PMB_NCOUNTERS * ((uint64_t)(max_shift) + 1) + 3)
int  pmb_run_chrom_noalloc(pmb_ctx *ctx, const pmb_chrom_args *a, int64_t *d_out, uint64_t out_int64);

/* Integer-pipe calibration on the device: runs a dependent-free loop of `what`
 * (0 = POPC, 1 = LOP3, 2 = POPC+LOP3 interleaved, 3 = SHF, 4 = IADD3, 5 = POPC+IMAD)
 * and returns executed thread-level ops per second. */
int  pmb_pipe_probe(pmb_ctx *ctx, int what, double *ops_per_sec, float *ms);

#ifdef __cplusplus
}
#endif
#endif /* PYMASC_B200_H */
