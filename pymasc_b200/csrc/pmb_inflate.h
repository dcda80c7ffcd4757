// pmb_inflate.h -- a raw-DEFLATE (RFC 1951) decoder for BGZF blocks, written for throughput:
// 64-bit bit buffer refilled eight bytes at a time, two-level decode tables indexed by the next
// bits of the stream (11 bits for literals/lengths, 8 for distances; longer codes go through
// fixed-size second-level tables), length/distance base and extra-bit count packed in the table
// entry, a main loop that runs without bounds checks while input and output have room to spare
// and takes up to three literals per refill (BAM payloads -- names, packed bases, qualities --
// are mostly literals; the ends of the block go through a checked loop, one symbol at a time),
// word-wide match copies.  (Tried and dropped: first-level slots holding two literals -- the
// longer table build and the variable advance cost more than the shorter dependency chain won.)
//
// It is an accelerator, not the authority: the caller knows the exact inflated size (ISIZE) and
// the CRC-32 of every BGZF block, checks both, and hands any block this decoder refuses (or gets
// wrong) to zlib.  So every branch below may simply `return false` on anything unexpected.
// Writes never go past out + out_len (neighbouring blocks are being written by other threads).
#ifndef PMB_INFLATE_H
#define PMB_INFLATE_H

#include <stdint.h>
#include <string.h>

namespace pmb_inflate {

constexpr int LL_BITS = 11;      // first-level index width, literal/length codes
constexpr int D_BITS = 8;        // first-level index width, distance codes
constexpr int MAX_LEN = 15;      // longest Huffman code of DEFLATE
// entry: value << 16 | extra << 8 | kind << 4 | code bits consumed at this level
//   LITERAL   value = the byte
//   BASE      value = base length / distance, extra = number of extra bits
//   SUBTABLE  value = offset of the second-level table, extra = its index width
enum Kind : uint32_t { LITERAL = 0, BASE = 1, END_OF_BLOCK = 2, SUBTABLE = 3, INVALID = 4 };
constexpr uint32_t entry(uint32_t value, uint32_t extra, uint32_t kind, uint32_t bits)
{
    return value << 16 | extra << 8 | kind << 4 | bits;
}
// worst case: every code longer than the first level gets a second-level table of its own
constexpr int LL_TABLE = (1 << LL_BITS) + 288 * (1 << (MAX_LEN - LL_BITS));
constexpr int D_TABLE = (1 << D_BITS) + 32 * (1 << (MAX_LEN - D_BITS));

struct Tables {
    uint32_t ll[LL_TABLE];
    uint32_t d[D_TABLE];
};

inline uint32_t reverse_bits(uint32_t code, int len)
{
    uint32_t r = 0;
    for (int i = 0; i < len; i++) r |= ((code >> i) & 1u) << (len - 1 - i);
    return r;
}

// symbol -> entry payload (value, extra, kind) of the two alphabets
inline uint32_t ll_payload(int sym)
{
    static const uint16_t base[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59,
                                      67, 83, 99, 115, 131, 163, 195, 227, 258};
    static const uint8_t extra[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
    if (sym < 256) return entry((uint32_t)sym, 0, LITERAL, 0);
    if (sym == 256) return entry(0, 0, END_OF_BLOCK, 0);
    if (sym <= 285) return entry(base[sym - 257], extra[sym - 257], BASE, 0);
    return entry(0, 0, INVALID, 0);
}
inline uint32_t d_payload(int sym)
{
    static const uint16_t base[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769,
                                      1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
    static const uint8_t extra[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};
    if (sym < 30) return entry(base[sym], extra[sym], BASE, 0);
    return entry(0, 0, INVALID, 0);
}

// Canonical Huffman code (RFC 1951 3.2.2) from code lengths -> two-level table.  Codes no longer
// than `first` bits fill every first-level slot whose low bits equal the (bit-reversed) code;
// longer codes share a second-level table per `first`-bit prefix, all of 2^(longest-first)
// slots.  Returns false for an over-subscribed code; an incomplete one leaves INVALID slots
// (legal for a single distance code, harmless otherwise: hitting one fails the block).
template <class Payload>
inline bool build_table(const uint8_t* lens, int n_syms, int first, uint32_t* table, int capacity, Payload payload)
{
    int count[MAX_LEN + 1] = {0};
    for (int s = 0; s < n_syms; s++) count[lens[s]]++;
    count[0] = 0;
    int longest = 0;
    for (int l = 1; l <= MAX_LEN; l++)
        if (count[l]) longest = l;
    uint32_t next_code[MAX_LEN + 2] = {0};
    uint32_t code = 0;
    long long space = 1;
    for (int l = 1; l <= MAX_LEN; l++) {
        code = (code + (uint32_t)count[l - 1]) << 1;
        next_code[l] = code;
        space = space * 2 - count[l];
        if (space < 0) return false;
    }
    const int size1 = 1 << first;
    for (int i = 0; i < size1; i++) table[i] = entry(0, 0, INVALID, 1);
    const int sub_bits = longest > first ? longest - first : 0;
    int used = size1;
    for (int s = 0; s < n_syms; s++) {
        const int l = lens[s];
        if (!l) continue;
        const uint32_t rev = reverse_bits(next_code[l]++, l);
        if (l <= first) {
            const uint32_t e = payload(s) | (uint32_t)l;
            for (uint32_t i = rev; i < (uint32_t)size1; i += 1u << l) table[i] = e;
            continue;
        }
        const uint32_t prefix = rev & (uint32_t)(size1 - 1);
        if (((table[prefix] >> 4) & 7u) != SUBTABLE) {
            if (used + (1 << sub_bits) > capacity) return false;
            table[prefix] = entry((uint32_t)used, (uint32_t)sub_bits, SUBTABLE, (uint32_t)first);
            for (int i = 0; i < (1 << sub_bits); i++) table[used + i] = entry(0, 0, INVALID, 1);
            used += 1 << sub_bits;
        }
        uint32_t* sub = table + (table[prefix] >> 16);
        const uint32_t e = payload(s) | (uint32_t)(l - first);
        for (uint32_t i = rev >> first; i < (1u << sub_bits); i += 1u << (l - first)) sub[i] = e;
    }
    return true;
}

struct BitReader {
    const uint8_t* in;
    const uint8_t* end;
    uint64_t buf = 0;
    int bits = 0;           // valid bits in buf (may go negative: more consumed than the input held)
    inline void refill()
    {
        if (end - in >= 8) {            // branch-free bulk refill: top up to at least 56 bits
            uint64_t w;
            memcpy(&w, in, 8);
            buf |= w << bits;
            in += (63 - bits) >> 3;
            bits |= 56;
        } else {
            while (bits <= 56 && in < end) {
                buf |= (uint64_t)*in++ << bits;
                bits += 8;
            }
        }
    }
    inline void refill_fast()       // the caller knows end - in >= 8 and 0 <= bits <= 63
    {
        uint64_t w;
        memcpy(&w, in, 8);
        buf |= w << bits;
        in += (63 - bits) >> 3;
        bits |= 56;
    }
    inline uint32_t peek(int n) const { return (uint32_t)buf & ((1u << n) - 1u); }
    inline void drop(int n) { buf >>= n; bits -= n; }
    inline uint32_t take(int n) { const uint32_t v = peek(n); drop(n); return v; }
};

inline const Tables& fixed_tables()
{
    static const Tables* t = [] {
        Tables* x = new Tables();
        uint8_t lens[288];
        for (int s = 0; s < 288; s++) lens[s] = s < 144 ? 8 : s < 256 ? 9 : s < 280 ? 7 : 8;
        build_table(lens, 288, LL_BITS, x->ll, LL_TABLE, ll_payload);
        uint8_t dl[32];
        for (int s = 0; s < 32; s++) dl[s] = 5;
        build_table(dl, 32, D_BITS, x->d, D_TABLE, d_payload);
        return x;
    }();
    return *t;
}

// dynamic block header (RFC 1951 3.2.7) -> tables
inline bool read_dynamic_tables(BitReader& br, Tables& t)
{
    br.refill();
    const int hlit = (int)br.take(5) + 257, hdist = (int)br.take(5) + 1, hclen = (int)br.take(4) + 4;
    if (hlit > 286 || hdist > 30) return false;
    static const uint8_t order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
    uint8_t cl[19] = {0};
    for (int i = 0; i < hclen; i++) {
        if (br.bits < 3) br.refill();
        cl[order[i]] = (uint8_t)br.take(3);
    }
    if (br.bits < 0) return false;
    uint32_t pre[1 << 7];
    auto pre_payload = [](int s) { return entry((uint32_t)s, 0, LITERAL, 0); };
    if (!build_table(cl, 19, 7, pre, 1 << 7, pre_payload)) return false;
    uint8_t lens[286 + 30 + 138];
    int n = 0;
    while (n < hlit + hdist) {
        br.refill();
        const uint32_t e = pre[br.peek(7)];
        if (((e >> 4) & 7u) != LITERAL) return false;
        br.drop((int)(e & 15u));
        const int sym = (int)(e >> 16);
        if (sym < 16) { lens[n++] = (uint8_t)sym; continue; }
        int rep;
        uint8_t v = 0;
        if (sym == 16) {
            if (n == 0) return false;
            v = lens[n - 1];
            rep = 3 + (int)br.take(2);
        } else if (sym == 17) rep = 3 + (int)br.take(3);
        else rep = 11 + (int)br.take(7);
        if (n + rep > hlit + hdist) return false;
        memset(lens + n, v, (size_t)rep);
        n += rep;
    }
    if (br.bits < 0 || lens[256] == 0) return false;
    return build_table(lens, hlit, LL_BITS, t.ll, LL_TABLE, ll_payload) &&
           build_table(lens + hlit, hdist, D_BITS, t.d, D_TABLE, d_payload);
}

// Inflate in[0, in_len) into exactly out[0, out_len).  false: not decoded (use zlib).
inline bool inflate(const uint8_t* in, size_t in_len, uint8_t* out, size_t out_len)
{
    BitReader br{in, in + in_len};
    uint8_t* const out0 = out;
    uint8_t* const out_end = out + out_len;
    Tables dyn;
    for (bool last = false; !last;) {
        br.refill();
        last = br.take(1) != 0;
        const uint32_t type = br.take(2);
        if (br.bits < 0) return false;
        if (type == 0) {                // stored: back to a byte boundary, LEN, ~LEN, bytes
            br.drop(br.bits & 7);
            br.refill();
            if (br.bits < 32) return false;
            const uint32_t len = br.take(16), nlen = br.take(16);
            if ((len ^ nlen) != 0xFFFFu) return false;
            // hand the bytes still in the bit buffer back to the input
            br.in -= br.bits >> 3;
            br.buf = 0;
            br.bits = 0;
            if ((size_t)(br.end - br.in) < len || (size_t)(out_end - out) < len) return false;
            memcpy(out, br.in, len);
            br.in += len;
            out += len;
            continue;
        }
        if (type == 3) return false;
        const Tables* t = &fixed_tables();
        if (type == 2) {
            if (!read_dynamic_tables(br, dyn)) return false;
            t = &dyn;
        }
        // length and distance of a match whose length code is `e` (already dropped); false = refuse
        auto copy_match = [&](uint32_t e, bool roomy) -> bool {
            const uint32_t len = (e >> 16) + br.take((int)((e >> 8) & 0xFFu));
            br.refill();
            uint32_t d = t->d[br.peek(D_BITS)];
            if ((d & 0x70u) == (SUBTABLE << 4)) {
                br.drop(D_BITS);
                d = t->d[(d >> 16) + br.peek((int)((d >> 8) & 0xFFu))];
            }
            br.drop((int)(d & 15u));
            if ((d & 0x70u) != (BASE << 4)) return false;
            const uint32_t dist = (d >> 16) + br.take((int)((d >> 8) & 0xFFu));
            if (br.bits < 0 || dist > (size_t)(out - out0) || len > (size_t)(out_end - out)) return false;
            const uint8_t* src = out - dist;
            uint8_t* const stop = out + len;
            if (roomy || (size_t)(out_end - out) >= len + 16) {     // copies may overshoot by < 16
                if (dist >= 16) {
                    do {
                        memcpy(out, src, 16);
                        out += 16;
                        src += 16;
                    } while (out < stop);
                } else if (dist >= 8) {
                    do {
                        memcpy(out, src, 8);
                        out += 8;
                        src += 8;
                    } while (out < stop);
                } else if (dist == 1) {
                    memset(out, *src, len);
                } else {
                    do *out++ = *src++; while (out < stop);
                }
                out = stop;
            } else {
                do *out++ = *src++; while (out < stop);
            }
            return true;
        };
        auto is_literal = [](uint32_t e) { return (e & 0x70u) == (LITERAL << 4); };
        constexpr size_t ROOM = 3 + 1 + 258 + 16 + 8;       // what one round of the fast loop may write, and more
        for (bool end_of_block = false; !end_of_block;) {
            // ---- fast: no bounds checks while >= 16 bytes of input and ROOM bytes of output are left
            while ((size_t)(br.end - br.in) >= 16 && (size_t)(out_end - out) >= ROOM) {
                br.refill_fast();                       // >= 56 bits: three first-level codes and a length fit
                uint32_t e = t->ll[br.peek(LL_BITS)];
                #define PMB_EMIT(e) do { *out++ = (uint8_t)((e) >> 16); br.drop((int)((e) & 15u)); } while (0)
                if (is_literal(e)) {
                    PMB_EMIT(e);
                    e = t->ll[br.peek(LL_BITS)];
                    if (is_literal(e)) {
                        PMB_EMIT(e);
                        e = t->ll[br.peek(LL_BITS)];
                        if (is_literal(e)) { PMB_EMIT(e); continue; }
                    }
                }
                if ((e & 0x70u) == (SUBTABLE << 4)) {
                    br.drop(LL_BITS);
                    e = t->ll[(e >> 16) + br.peek((int)((e >> 8) & 0xFFu))];
                    if (is_literal(e)) { PMB_EMIT(e); continue; }
                }
                #undef PMB_EMIT
                br.drop((int)(e & 15u));
                const uint32_t kind = (e >> 4) & 7u;
                if (kind == END_OF_BLOCK) { end_of_block = true; break; }
                if (kind != BASE || !copy_match(e, true)) return false;
            }
            if (end_of_block) break;
            // ---- checked: one symbol
            br.refill();
            uint32_t e = t->ll[br.peek(LL_BITS)];
            if ((e & 0x70u) == (SUBTABLE << 4)) {
                br.drop(LL_BITS);
                e = t->ll[(e >> 16) + br.peek((int)((e >> 8) & 0xFFu))];
            }
            br.drop((int)(e & 15u));
            const uint32_t kind = (e >> 4) & 7u;
            if (kind == LITERAL) {
                if (out == out_end) return false;
                *out++ = (uint8_t)(e >> 16);
                continue;
            }
            if (kind == END_OF_BLOCK) break;
            if (kind != BASE || !copy_match(e, false)) return false;
        }
        if (br.bits < 0) return false;
    }
    return out == out_end;
}

}  // namespace pmb_inflate
#endif
