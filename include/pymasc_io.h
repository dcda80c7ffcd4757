/*
 * pymasc_io.h -- C ABI of libpymasc_io.so: host-side BAM (BGZF) and BigWig decoding into the
 * packed arrays that libpymasc_b200.so consumes (include/pymasc_b200.h).  C++ behind a plain C
 * ABI: its own inflate and CRC-32 for BGZF blocks with zlib as the fallback, std::thread; no
 * CUDA in this library.  north_star keeps BAM/BigWig decoding on the host; the
 * reference does it through pysam and pyBigWig, neither of which exists in this image.
 *
 * Reference interfaces replaced (paths relative to the PyMaSC checkout):
 *   pmio_bam_load ............ the `for read in AlignmentFile(path)` loops of
 *                              PyMaSC/handler/calc.py:140-153 and handler/worker.py:126-132 with
 *                              the filter of handler/read.py:62-90 (read2, MAPQ, unmapped,
 *                              duplicate) and the coordinate rules of handler/read.py:132-149
 *                              (pos = reference_start + 1, readlen = infer_query_length()).
 *   pmio_bam_readlen_hist .... the length counter of estimate_readlen, core/readlen.pyx:130-160
 *                              (not unmapped, not duplicate, MAPQ >= criteria; read2 included).
 *   pmio_bw_open / _intervals  BigWigReader.__init__ / .fetch, reader/bigwig.pyx:114-177
 *                              (pyBigWig.open, .chroms(), .intervals(chrom, 0, size), value filter).
 *   pmio_bw_prefetch ......... the same decode for every chromosome at once, on a thread pool
 *                              (the reference fetches per chromosome and per input file).
 *   pmio_inflate_raw[2], pmio_crc32 ... test hooks of the block decoder (no reference counterpart:
 *                              pysam's htslib does this inside).
 *
 * Every function returns 0 or a negative pmio_status; text in pmio_last_error() (thread-local).
 * Arrays returned by the getters are owned by the handle and live until it is freed (BigWig:
 * until the next pmio_bw_intervals / pmio_bw_prefetch call on the same handle).
 */
#ifndef PYMASC_IO_H
#define PYMASC_IO_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    PMIO_OK = 0,
    PMIO_ERR_IO = -1,       /* open / read failed                                   */
    PMIO_ERR_FORMAT = -2,   /* not a BAM / BigWig file, or a corrupt one            */
    PMIO_ERR_ARG = -3,
    PMIO_ERR_NOMEM = -4,
    PMIO_ERR_KEY = -5       /* chromosome not in the file (KeyError on the Python side) */
} pmio_status;

const char *pmio_last_error(void);

/* One raw-DEFLATE stream (RFC 1951, no zlib/gzip wrapper) into exactly out_len bytes, by the
 * library's own decoder (use_zlib = 0; pymasc_b200/csrc/pmb_inflate.h -- what BGZF blocks go
 * through first) or by zlib (use_zlib = 1; the fallback for anything the former refuses).
 * PMIO_ERR_FORMAT when the stream does not decode to that size. */
int pmio_inflate_raw(const uint8_t *in, uint64_t in_len, uint8_t *out, uint64_t out_len, int use_zlib);
/* Two independent raw-DEFLATE streams decoded together by the library's own decoder (their
 * decode loops take turns; what pairs of BGZF blocks go through).  ok[i] = 1 when stream i decoded
 * to exactly out_len[i] bytes.  Always PMIO_OK unless an argument is null. */
int pmio_inflate_raw2(const uint8_t *const in[2], const uint64_t in_len[2], uint8_t *const out[2],
                      const uint64_t out_len[2], int ok[2]);

/* CRC-32 (gzip) of a buffer, by carry-less multiplication where the CPU has it (use_zlib = 0;
 * pymasc_b200/csrc/pmb_crc32.h -- what checks every BGZF block) or by zlib's crc32 (use_zlib = 1). */
uint32_t pmio_crc32(const uint8_t *data, uint64_t len, int use_zlib);

/* ------------------------------------------------------------------ BAM */
typedef struct pmio_bam pmio_bam;

typedef struct {
    uint64_t n_records;     /* alignment records in the file                               */
    uint64_t n_with_ref;    /* records with a reference (readlen.pyx: `nreads`)           */
    uint64_t n_unmapped;    /* of those, flag 0x4                                          */
    uint64_t n_paired;      /* of those, flag 0x1                                          */
    uint64_t n_read2;       /* of those, paired and flag 0x80                              */
    uint64_t n_accepted;    /* records handed to the calculator (handler/read.py rules)   */
    uint32_t sorted;        /* 1 if the accepted reads are coordinate sorted: no reference */
                            /* revisited, positions non-decreasing (mscc.pyx:351-366)     */
    uint32_t reparsed_shares; /* decoder statistic: shares whose guessed first record (found   */
                            /* from the bytes at a BGZF block boundary) was wrong and were    */
                            /* filtered again from the true boundary                          */
} pmio_bam_stats;

/* Decode the whole file: BGZF blocks are inflated on `threads` threads (0 = all cores), records
 * are filtered (skip read2, MAPQ < mapq_criteria, unmapped, duplicate, no reference, no CIGAR)
 * and packed per reference in file order.
 * Environment switches, read by the library (measurement and test aids, not configuration):
 *   PMIO_TIMING=1        seconds per stage on stderr (also pmio_bw_prefetch)
 *   PMIO_ZLIB_ONLY=1     zlib's inflate / crc32 / uncompress instead of the library's own (A/B;
 *                        read once per process)
 *   PMIO_BATCH_BLOCKS=n  BGZF blocks per batch (default 128 per thread, 512..4096)
 *   PMIO_SHARE_BYTES=n   inflated bytes per share (default 512 KiB; tests shrink it) */
int  pmio_bam_load(const char *path, int mapq_criteria, int threads, pmio_bam **out);
void pmio_bam_free(pmio_bam *b);
int  pmio_bam_stats_get(const pmio_bam *b, pmio_bam_stats *s);
uint32_t    pmio_bam_n_refs(const pmio_bam *b);
const char *pmio_bam_ref_name(const pmio_bam *b, uint32_t i);
uint64_t    pmio_bam_ref_len(const pmio_bam *b, uint32_t i);
/* Accepted reads of reference i: 1-based leftmost base, query length (CIGAR M/I/S/=/X), strand
 * (1 = reverse).  The arrays are assembled on the first request that asks for them (pass NULL
 * pointers to get the count only). */
int  pmio_bam_reads(pmio_bam *b, uint32_t i, const uint32_t **pos, const uint32_t **qlen,
                    const uint8_t **is_reverse, uint64_t *n);
/* The same reads reduced by the calculator's feed rules (mscc.pyx:370-418), for a reference
 * whose own reads are coordinate sorted (pmio_bam_ref_sorted; PMIO_ERR_ARG otherwise): sorted unique forward 5' positions (consecutive
 * repeats dropped) and sorted unique reverse 5' ends pos + qlen - 1, each with the summed query
 * length of the reads that set a bit first; first/last leftmost position fed. */
int  pmio_bam_packed(const pmio_bam *b, uint32_t i, const uint32_t **fwd, uint64_t *n_fwd,
                     uint64_t *fwd_len_sum, const uint32_t **rev, uint64_t *n_rev,
                     uint64_t *rev_len_sum, uint32_t *first_pos, uint32_t *last_pos);
/* Coordinate order as the reference sees it: it checks order only on reads that survive the
 * chromosome filter (handler/calc.py:143-144 -> _check_pos, mscc.pyx:351-366), so the file-wide
 * `sorted` flag of pmio_bam_stats is not enough.
 *   pmio_bam_ref_sorted: 1 when the accepted reads of reference i have non-decreasing positions in
 *       file order (whatever other references lie between them), else 0.
 *   pmio_bam_ref_order:  the reference ids in the order their reads appear in the file, consecutive
 *       repeats merged (a reference listed twice was left and revisited); *overflow = 1 when the
 *       list was cut at 4 M entries (a hopelessly unsorted file). */
int  pmio_bam_ref_sorted(const pmio_bam *b, uint32_t i);
int  pmio_bam_ref_order(const pmio_bam *b, const int32_t **refs, uint64_t *n, int *overflow);
/* Query-length histogram for the read-length estimator (records without a CIGAR have no query
 * length and are not counted), in the order the lengths first occur in the file -- the order the
 * reference's MODE estimator breaks ties by (core/readlen.pyx:81). */
int  pmio_bam_readlen_hist(const pmio_bam *b, const uint32_t **lengths, const uint64_t **counts,
                           uint32_t *n);

/* ------------------------------------------------------------------ BigWig */
typedef struct pmio_bw pmio_bw;

int  pmio_bw_open(const char *path, pmio_bw **out);
void pmio_bw_close(pmio_bw *w);
uint32_t    pmio_bw_n_chroms(const pmio_bw *w);
const char *pmio_bw_chrom_name(const pmio_bw *w, uint32_t i);
uint64_t    pmio_bw_chrom_len(const pmio_bw *w, uint32_t i);
/* All intervals of `chrom` with value >= threshold (every interval when threshold <= 0, as
 * reader/bigwig.pyx:171-175 does), in file order: 0-based half-open begins/ends and the
 * float32 values.  PMIO_ERR_KEY when the file has no such chromosome. */
int  pmio_bw_intervals(pmio_bw *w, const char *chrom, float threshold, const uint32_t **begins,
                       const uint32_t **ends, const float **values, uint64_t *n);
/* Decodes EVERY chromosome at `threshold` on `threads` threads (0 = all cores) and keeps the
 * arrays in the handle: pmio_bw_intervals calls with the same threshold then return them without
 * decoding (pointers valid until pmio_bw_close or the next prefetch).  What a run over all
 * chromosomes (handler/mappability.py, the calculator's per-chromosome fetch) wants. */
int  pmio_bw_prefetch(pmio_bw *w, float threshold, int threads);

#ifdef __cplusplus
}
#endif
#endif /* PYMASC_IO_H */
