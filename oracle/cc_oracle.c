/*
 * cc_oracle.c -- CPU restatement of PyMaSC's bit-array cross-correlation path.
 *
 * TEST INFRASTRUCTURE ONLY.  This file is the parity checker for the CUDA
 * path in pymasc_b200/csrc.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load it.  The product
 * (pymasc_b200) never links, imports or falls back to anything in oracle/.
 *
 * Parity status: PINNED.  tests/test_oracle_golden.py checks this restatement
 * against the reference's own golden tables (tests/golden/ENCFF000RMB-test_*.tab,
 * tests/data/hg19_36mer-test_mappability.json) through the committed fixtures in
 * tests/golden/, and -- when /root/reference is present -- against the compiled
 * reference calculator built into oracle/_ref by oracle/build_ref.py.
 *
 * What is restated (reference file:line, relative to the PyMaSC checkout):
 *   bit layout, 64-bit LSB-first words ....... external/BitArray/bit_macros.h:68-69
 *   set one bit ............................... external/BitArray/bit_array.c:668
 *   fill a region [start, start+len) .......... external/BitArray/bit_array.c:423-479
 *   popcount of an array ...................... external/BitArray/bit_array.c:962-978
 *   unaligned 64-bit fetch (zero past end) .... external/BitArray/bit_array.c:311-330
 *   shift toward index 0 / toward the top ..... external/BitArray/bit_array.c:1953-2011
 *   AND + popcount without a temporary ........ PyMaSC/core/bitarray/bitarray.pyx:109-133
 *   per-shift loop ............................ PyMaSC/core/bitarray/mscc.pyx:288-317
 *   read-less chromosome mappable length ...... PyMaSC/core/bitarray/mscc.pyx:207-215
 *
 * The reference shifts whole arrays by one bit per iteration.  After k
 * iterations its reverse array holds R[j+k] at index j (zero filled), and its
 * "reverse mappability" holds M[j + read_len-1-k] at index j (mscc.pyx:283-284
 * initial rshift by read_len-1, then one lshift per iteration refilled from the
 * saved low bits, :307-310).  Instead of moving memory this file fetches the
 * unaligned word at the equivalent bit offset; the integers are identical.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORC_OK 0
#define ORC_ERR_RANGE (-2)

static inline uint64_t nwords64(uint64_t nbits) { return (nbits + 63) >> 6; }

/* bits [off, off+64) of an nbits-long array; zero outside [0, nbits). */
static inline uint64_t fetch64(const uint64_t *w, uint64_t nbits, int64_t off)
{
    if (off <= -64 || off >= (int64_t)nbits) return 0;
    uint64_t nw = nwords64(nbits);
    if (off < 0) {
        /* only the top (64+off) bits come from word 0 */
        return w[0] << (uint64_t)(-off);
    }
    uint64_t wi = (uint64_t)off >> 6, bo = (uint64_t)off & 63;
    uint64_t lo = w[wi] >> bo;
    if (bo && wi + 1 < nw) lo |= w[wi + 1] << (64 - bo);
    return lo;
}

/* bit_array_set_bit for every position (bit_array.c:668); asserts become an
 * error return (the reference aborts). */
int orc_set_bits(uint64_t *w, uint64_t nbits, const uint32_t *pos, uint64_t n)
{
    for (uint64_t i = 0; i < n; i++) {
        uint64_t p = pos[i];
        if (p >= nbits) return ORC_ERR_RANGE;
        w[p >> 6] |= (uint64_t)1 << (p & 63);
    }
    return ORC_OK;
}

/* mscc.pyx:343-344 + bitarray.pyx:88-95: interval (b, e) sets bits b+1 .. e. */
int orc_fill_intervals(uint64_t *w, uint64_t nbits, const uint32_t *begins,
                       const uint32_t *ends, uint64_t n)
{
    for (uint64_t i = 0; i < n; i++) {
        uint64_t s = (uint64_t)begins[i] + 1, e = ends[i]; /* inclusive */
        if (e < s) continue;                               /* length 0 */
        if (e >= nbits) return ORC_ERR_RANGE;
        uint64_t fw = s >> 6, lw = e >> 6;
        uint64_t fm = ~(uint64_t)0 << (s & 63);
        uint64_t lm = ((e & 63) == 63) ? ~(uint64_t)0 : (((uint64_t)1 << ((e & 63) + 1)) - 1);
        if (fw == lw) {
            w[fw] |= fm & lm;
        } else {
            w[fw] |= fm;
            for (uint64_t j = fw + 1; j < lw; j++) w[j] = ~(uint64_t)0;
            w[lw] |= lm;
        }
    }
    return ORC_OK;
}

uint64_t orc_count(const uint64_t *w, uint64_t nbits)
{
    uint64_t nw = nwords64(nbits), c = 0;
    for (uint64_t i = 0; i < nw; i++) c += (uint64_t)__builtin_popcountll(w[i]);
    return c;
}

/*
 * The per-shift loop of _calc_correlation (mscc.pyx:288-317), for shifts
 * k = 0 .. max_shift, restricted to positions j in [lo, hi) (lo = 0,
 * hi = nbits for a whole chromosome; the restriction is what a position
 * shard computes).
 *
 *   ncc[k]  = sum_j F[j] R[j+k]                         (mscc.pyx:314)
 *   dmp[k]  = sum_j M[j] M[j+rl-1-k]                    (:291-298, count of D)
 *   mfwd[k] = sum_j F[j] D_k[j]                         (:300,:303)
 *   mrev[k] = sum_j R[j+k] D_k[j]                       (:301,:304)
 *   mscc[k] = sum_j F[j] R[j+k] D_k[j]                  (:305)
 *
 * F, R, M are nbits long; M may be NULL (NCC only); ncc may be NULL
 * (skip_ncc).  Output arrays hold max_shift+1 int64 each.
 */
int orc_cc_counts(const uint64_t *F, const uint64_t *R, const uint64_t *M,
                  uint64_t nbits, int64_t read_len, int64_t max_shift,
                  uint64_t lo, uint64_t hi,
                  int64_t *ncc, int64_t *dmp, int64_t *mfwd, int64_t *mrev,
                  int64_t *mscc)
{
    if (hi > nbits) hi = nbits;
    if (lo > hi) lo = hi;
    uint64_t w0 = lo >> 6, w1 = (hi + 63) >> 6; /* words touched */
#pragma omp parallel for schedule(dynamic, 4)
    for (int64_t k = 0; k <= max_shift; k++) {
        int64_t moff = read_len - 1 - k;
        uint64_t c_ncc = 0, c_d = 0, c_f = 0, c_r = 0, c_m = 0;
        for (uint64_t i = w0; i < w1; i++) {
            uint64_t mask = ~(uint64_t)0;
            if (i == w0 && (lo & 63)) mask &= ~(uint64_t)0 << (lo & 63);
            if (i == w1 - 1 && (hi & 63)) mask &= (((uint64_t)1 << (hi & 63)) - 1);
            uint64_t f = F[i] & mask;
            uint64_t m = M ? (M[i] & mask) : 0;
            if (!f && !m) continue; /* every term carries F[j] or M[j] */
            uint64_t r = fetch64(R, nbits, (int64_t)(i << 6) + k);
            if (ncc) c_ncc += (uint64_t)__builtin_popcountll(f & r);
            if (M) {
                uint64_t d = m & fetch64(M, nbits, (int64_t)(i << 6) + moff);
                c_d += (uint64_t)__builtin_popcountll(d);
                c_f += (uint64_t)__builtin_popcountll(f & d);
                c_r += (uint64_t)__builtin_popcountll(r & d);
                c_m += (uint64_t)__builtin_popcountll(f & r & d);
            }
        }
        if (ncc) ncc[k] = (int64_t)c_ncc;
        if (M) {
            dmp[k] = (int64_t)c_d; mfwd[k] = (int64_t)c_f;
            mrev[k] = (int64_t)c_r; mscc[k] = (int64_t)c_m;
        }
    }
    return ORC_OK;
}

/* _fill_result for a read-less chromosome (mscc.pyx:207-215):
 * out[i] = sum_j M[j] M[j+i], i = 0 .. max_shift. */
int orc_mappable_autocorr(const uint64_t *M, uint64_t nbits, int64_t max_shift,
                          int64_t *out)
{
    uint64_t nw = nwords64(nbits);
#pragma omp parallel for schedule(dynamic, 4)
    for (int64_t k = 0; k <= max_shift; k++) {
        uint64_t c = 0;
        for (uint64_t i = 0; i < nw; i++) {
            uint64_t m = M[i];
            if (!m) continue;
            c += (uint64_t)__builtin_popcountll(m & fetch64(M, nbits, (int64_t)(i << 6) + k));
        }
        out[k] = (int64_t)c;
    }
    return ORC_OK;
}

/* popcount restricted to [lo, hi): the per-shard strand totals. */
uint64_t orc_count_range(const uint64_t *w, uint64_t nbits, uint64_t lo, uint64_t hi)
{
    if (hi > nbits) hi = nbits;
    uint64_t c = 0;
    for (uint64_t p = lo; p < hi;) {
        if ((p & 63) == 0 && p + 64 <= hi) {
            c += (uint64_t)__builtin_popcountll(w[p >> 6]);
            p += 64;
        } else {
            c += (w[p >> 6] >> (p & 63)) & 1;
            p++;
        }
    }
    return c;
}
