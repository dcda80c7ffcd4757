// pmb_crc32.h -- CRC-32 (the gzip / BGZF one: polynomial 0x04C11DB7, bit-reflected, initial value
// and final XOR 0xFFFFFFFF) with carry-less multiplication, for the 64 KiB blocks of a BAM file;
// zlib's table-driven crc32() costs as much as inflating a well-compressed block.
//
// The message is a polynomial over GF(2); the first byte holds the highest degrees, bit 0 of a
// byte the highest of that byte (the "reflected" convention), so the low half of a little-endian
// 128-bit load is the high-degree half.  A 128-bit accumulator A that stands D bits before the
// data it is added to is moved there by
//     A' = A.lo * (x^(D+32) mod P)  ^  A.hi * (x^(D-32) mod P)          (carry-less products)
// which is congruent to A * x^D modulo P; the two constants below are those residues for D = 512
// (four accumulators striding 64 bytes) and D = 128, bit-reflected and shifted left once (a
// reflected 64x64 product comes out one bit low in the 128-bit result).  They were computed with
//     r = 1; repeat k times: r <<= 1; if r >> 32: r ^= 0x104C11DB7        (x^k mod P)
// and are checked, with everything else here, against zlib on every length (tests/test_io_native.py).
// The last step is not the usual Barrett reduction: once the data is folded into 16 bytes, the
// CRC of the message equals the CRC of those 16 bytes followed by the < 16 bytes left over, and
// that short stretch goes through zlib's crc32().  The initial value 0xFFFFFFFF is the same as
// XOR-ing the first four message bytes with FF.
#ifndef PMB_CRC32_H
#define PMB_CRC32_H

#include <stdint.h>
#include <string.h>
#include <zlib.h>

#if defined(__x86_64__) && defined(__GNUC__)
#include <immintrin.h>
#define PMB_CRC32_CLMUL 1
#endif

namespace pmb_crc32 {

#ifdef PMB_CRC32_CLMUL
// A moved `D` bits up (see above): k = {residue for the low half, residue for the high half}
__attribute__((target("pclmul,sse4.1"))) static inline __m128i shift_up(__m128i a, __m128i k)
{
    return _mm_xor_si128(_mm_clmulepi64_si128(a, k, 0x00), _mm_clmulepi64_si128(a, k, 0x11));
}

// n >= 64
__attribute__((target("pclmul,sse4.1"))) static inline uint32_t clmul(const uint8_t* p, size_t n)
{
    const __m128i by512 = _mm_set_epi64x(0x1c6e41596, 0x154442bd4);     // x^(512-32), x^(512+32)
    const __m128i by128 = _mm_set_epi64x(0x0ccaa009e, 0x1751997d0);     // x^(128-32), x^(128+32)
    __m128i a0 = _mm_loadu_si128((const __m128i*)(p + 0));
    __m128i a1 = _mm_loadu_si128((const __m128i*)(p + 16));
    __m128i a2 = _mm_loadu_si128((const __m128i*)(p + 32));
    __m128i a3 = _mm_loadu_si128((const __m128i*)(p + 48));
    a0 = _mm_xor_si128(a0, _mm_cvtsi32_si128((int)0xFFFFFFFFu));
    p += 64;
    n -= 64;
    for (; n >= 64; p += 64, n -= 64) {
        a0 = _mm_xor_si128(shift_up(a0, by512), _mm_loadu_si128((const __m128i*)(p + 0)));
        a1 = _mm_xor_si128(shift_up(a1, by512), _mm_loadu_si128((const __m128i*)(p + 16)));
        a2 = _mm_xor_si128(shift_up(a2, by512), _mm_loadu_si128((const __m128i*)(p + 32)));
        a3 = _mm_xor_si128(shift_up(a3, by512), _mm_loadu_si128((const __m128i*)(p + 48)));
    }
    __m128i a = _mm_xor_si128(shift_up(a0, by128), a1);
    a = _mm_xor_si128(shift_up(a, by128), a2);
    a = _mm_xor_si128(shift_up(a, by128), a3);
    for (; n >= 16; p += 16, n -= 16) a = _mm_xor_si128(shift_up(a, by128), _mm_loadu_si128((const __m128i*)p));
    uint8_t folded[16];
    _mm_storeu_si128((__m128i*)folded, a);
    // zlib's crc32(c, ...) continues from register ~c: starting it at 0xFFFFFFFF starts the register at 0
    uLong c = crc32(0xFFFFFFFFul, folded, 16);
    if (n) c = crc32(c, p, (uInt)n);
    return (uint32_t)c;
}

static inline bool have_clmul()
{
    static const bool yes = __builtin_cpu_supports("pclmul") && __builtin_cpu_supports("sse4.1");
    return yes;
}
#endif

// CRC-32 of p[0, n); use_clmul 0 forces zlib (tests, A/B)
static inline uint32_t compute(const uint8_t* p, size_t n, bool use_clmul = true)
{
#ifdef PMB_CRC32_CLMUL
    if (use_clmul && n >= 64 && have_clmul()) return clmul(p, n);
#endif
    uLong c = crc32(0L, Z_NULL, 0);
    while (n) {                     // uInt lengths
        const size_t step = n < (1u << 30) ? n : (1u << 30);
        c = crc32(c, p, (uInt)step);
        p += step;
        n -= step;
    }
    return (uint32_t)c;
}

}  // namespace pmb_crc32
#endif
