// pmb_inflate.h -- a raw-DEFLATE (RFC 1951) decoder for BGZF blocks, written for throughput:
// 64-bit bit buffer refilled eight bytes at a time, two-level decode tables indexed by the next
// bits of the stream (11 bits for literals/lengths, 8 for distances; longer codes go through
// fixed-size second-level tables), length/distance base and extra-bit count packed in the table
// entry, a main loop that runs without bounds checks while input and output have room to spare
// and takes up to three literals per refill (BAM payloads -- names, packed bases, qualities --
// are mostly literals; the ends of the block go through a checked loop, one symbol at a time),
// word-wide match copies; two streams can be decoded together (inflate2).  (Tried and dropped: first-level slots holding two literals -- the
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
//
// The symbols are taken in code order (by length, then by symbol: a counting sort), so that
//  * the bit-reversed code word is kept incrementally -- +1 on a mirrored number is "clear the
//    set bits from the top, set the first clear one" -- and stays the same number when the length
//    grows (the zeros appended to the code go to the top of its mirror image);
//  * the first level is built by doubling: a code of l bits is stored ONCE in the table of 2^l
//    slots that exists when its length is reached, and every step to the next length copies the
//    table onto its upper half (a slot's index gains a high bit that the code does not look at).
//    No per-symbol fill loops, whose trip counts the branch predictor cannot learn;
//  * codes longer than `first` bits arrive grouped by their first-level prefix, so a second-level
//    table is opened whenever the prefix changes.
template <class Payload>
inline bool build_table(const uint8_t* lens, int n_syms, int first, uint32_t* table, int capacity, Payload payload)
{
    int count[MAX_LEN + 1] = {0};
    for (int s = 0; s < n_syms; s++) count[lens[s]]++;
    count[0] = 0;
    int longest = 0, start[MAX_LEN + 2];
    long long space = 1;
    start[1] = 0;
    for (int l = 1; l <= MAX_LEN; l++) {
        if (count[l]) longest = l;
        start[l + 1] = start[l] + count[l];
        space = space * 2 - count[l];
        if (space < 0) return false;
    }
    uint16_t sorted[288];
    {
        int at[MAX_LEN + 1];
        for (int l = 1; l <= MAX_LEN; l++) at[l] = start[l];
        for (int s = 0; s < n_syms; s++)
            if (lens[s]) sorted[at[lens[s]]++] = (uint16_t)s;
    }
    const bool complete = space == 0;           // then every slot gets written below
    const int size1 = 1 << first;
    if (!complete)
        for (int i = 0; i < size1; i++) table[i] = entry(0, 0, INVALID, 1);
    uint32_t rev = 0;                           // mirror image of the next code word
    auto next_code = [&rev](int len) {
        const uint32_t all = (1u << len) - 1u;
        if (rev == all) { rev = 0; return; }    // that was the last code of a complete code
        const uint32_t bit = 1u << (31 - __builtin_clz(rev ^ all));
        rev = (rev & (bit - 1u)) | bit;
    };
    int k = 0, size = 1;
    for (int len = 1; len <= first; len++) {
        memcpy(table + size, table, (size_t)size * sizeof(uint32_t));
        size *= 2;
        for (int c = 0; c < count[len]; c++, k++) {
            table[rev] = payload((int)sorted[k]) | (uint32_t)len;
            next_code(len);
        }
    }
    const int sub_bits = longest > first ? longest - first : 0;
    int used = size1;
    uint32_t prefix_now = ~0u;
    uint32_t* sub = nullptr;
    for (int len = first + 1; len <= longest; len++) {
        for (int c = 0; c < count[len]; c++, k++) {
            const uint32_t prefix = rev & (uint32_t)(size1 - 1);
            if (prefix != prefix_now) {
                if (used + (1 << sub_bits) > capacity) return false;
                table[prefix] = entry((uint32_t)used, (uint32_t)sub_bits, SUBTABLE, (uint32_t)first);
                sub = table + used;
                if (!complete)
                    for (int i = 0; i < (1 << sub_bits); i++) sub[i] = entry(0, 0, INVALID, 1);
                used += 1 << sub_bits;
                prefix_now = prefix;
            }
            const uint32_t e = payload((int)sorted[k]) | (uint32_t)(len - first);
            for (uint32_t i = rev >> first; i < (1u << sub_bits); i += 1u << (len - first)) sub[i] = e;
            next_code(len);
        }
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
        if (br.bits < 14) br.refill();          // a code of <= 7 bits and <= 7 extra bits
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

// (inflate / inflate2 / inflate_upto are always inlined: the caller decides the instruction set,
// pmb_io.cpp clones its wrappers for BMI2, whose flag-free variable shifts are worth ~9 % here.)
// ---- the decoder.  State of one stream; `prepare` walks block headers, stored blocks and the
// checked one-symbol loop until the stream stands inside a Huffman block with room for the fast
// loop (or is finished / refused); `fast_round` is one refill's worth of that loop.  Two streams
// can take turns in it (inflate2): the decode of a symbol is one dependency chain -- table load,
// shift of the bit buffer by the code length just loaded, mask, next load -- and two independent
// chains overlap in the out-of-order core.
#define PMB_INLINE inline __attribute__((always_inline))

// what the fast loop keeps in registers
struct Hot {
    uint64_t buf;
    int bits;
    const uint8_t* in;
    uint8_t* out;
};

constexpr size_t FAST_IN = 16;                      // input bytes one round may read past `in`
constexpr size_t FAST_OUT = 3 + 1 + 258 + 16 + 8;   // output bytes one round may write, and more

enum Round { MORE = 0, BLOCK_END = 1, REFUSED = 2 };

// One round: up to three literals, or a literal from a second-level table, or a match, or the
// end of the block.  No bounds checks: the caller guarantees FAST_IN / FAST_OUT.
static PMB_INLINE Round fast_round(Hot& h, const Tables* t, const uint8_t* out0)
{
    auto refill = [&]() {
        uint64_t w;
        memcpy(&w, h.in, 8);
        h.buf |= w << h.bits;
        h.in += (63 - h.bits) >> 3;
        h.bits |= 56;
    };
    auto drop = [&](uint32_t n) { h.buf >>= n; h.bits -= (int)n; };
    auto take = [&](uint32_t n) { const uint32_t v = (uint32_t)h.buf & ((1u << n) - 1u); drop(n); return v; };
    auto is_literal = [](uint32_t e) { return (e & 0x70u) == (LITERAL << 4); };
    constexpr uint32_t LL_MASK = (1u << LL_BITS) - 1u, D_MASK = (1u << D_BITS) - 1u;
    refill();                                       // >= 56 bits: three first-level codes and a length fit
    uint32_t e = t->ll[(uint32_t)h.buf & LL_MASK];
    if (is_literal(e)) {
        *h.out++ = (uint8_t)(e >> 16); drop(e & 15u);
        e = t->ll[(uint32_t)h.buf & LL_MASK];
        if (is_literal(e)) {
            *h.out++ = (uint8_t)(e >> 16); drop(e & 15u);
            e = t->ll[(uint32_t)h.buf & LL_MASK];
            if (is_literal(e)) { *h.out++ = (uint8_t)(e >> 16); drop(e & 15u); return MORE; }
        }
    }
    if ((e & 0x70u) == (SUBTABLE << 4)) {
        drop(LL_BITS);
        e = t->ll[(e >> 16) + ((uint32_t)h.buf & ((1u << ((e >> 8) & 0xFFu)) - 1u))];
        if (is_literal(e)) { *h.out++ = (uint8_t)(e >> 16); drop(e & 15u); return MORE; }
    }
    drop(e & 15u);
    const uint32_t kind = (e >> 4) & 7u;
    if (kind == END_OF_BLOCK) return BLOCK_END;
    if (kind != BASE) return REFUSED;
    const uint32_t len = (e >> 16) + take((e >> 8) & 0xFFu);
    refill();
    uint32_t d = t->d[(uint32_t)h.buf & D_MASK];
    if ((d & 0x70u) == (SUBTABLE << 4)) {
        drop(D_BITS);
        d = t->d[(d >> 16) + ((uint32_t)h.buf & ((1u << ((d >> 8) & 0xFFu)) - 1u))];
    }
    drop(d & 15u);
    if ((d & 0x70u) != (BASE << 4)) return REFUSED;
    const uint32_t dist = (d >> 16) + take((d >> 8) & 0xFFu);
    if (dist > (size_t)(h.out - out0)) return REFUSED;
    const uint8_t* src = h.out - dist;
    uint8_t* const stop = h.out + len;
    if (dist >= 16) {                               // copies may overshoot by < 16
        do {
            memcpy(h.out, src, 16);
            h.out += 16;
            src += 16;
        } while (h.out < stop);
    } else if (dist >= 8) {
        do {
            memcpy(h.out, src, 8);
            h.out += 8;
            src += 8;
        } while (h.out < stop);
    } else if (dist == 1) {
        memset(h.out, *src, len);
    } else {
        do *h.out++ = *src++; while (h.out < stop);
    }
    h.out = stop;
    return MORE;
}

struct Decoder {
    BitReader br{nullptr, nullptr};
    uint8_t* out0 = nullptr;
    uint8_t* out = nullptr;
    uint8_t* out_end = nullptr;
    const Tables* t = nullptr;
    bool last = false;          // the block being decoded is the final one
    bool in_block = false;      // inside a Huffman block (t is its tables)
    bool done = false;          // the stream ended (exactly at out_end, unless `exact` is off)
    bool refused = false;
    bool exact = true;          // off: out_end is a capacity, the stream may end before it
    Tables dyn;

    void init(const uint8_t* in, size_t in_len, uint8_t* o, size_t out_len)
    {
        br = BitReader{in, in + in_len};
        out0 = out = o;
        out_end = o + out_len;
        last = in_block = done = refused = false;
        exact = true;
    }
    bool fast_ok() const { return (size_t)(br.end - br.in) >= FAST_IN && (size_t)(out_end - out) >= FAST_OUT; }
    Hot hot() const { return Hot{br.buf, br.bits, br.in, out}; }
    void keep(const Hot& h) { br.buf = h.buf; br.bits = h.bits; br.in = h.in; out = h.out; }
    bool live() const { return !done && !refused; }
    void end_block() { in_block = false; }

    // block header (RFC 1951 3.2.3); a stored block is copied here
    void open_block()
    {
        br.refill();
        last = br.take(1) != 0;
        const uint32_t type = br.take(2);
        if (br.bits < 0 || type == 3) { refused = true; return; }
        if (type == 0) {                // stored: back to a byte boundary, LEN, ~LEN, bytes
            br.drop(br.bits & 7);
            br.refill();
            if (br.bits < 32) { refused = true; return; }
            const uint32_t len = br.take(16), nlen = br.take(16);
            if ((len ^ nlen) != 0xFFFFu) { refused = true; return; }
            br.in -= br.bits >> 3;      // hand the bytes still in the bit buffer back to the input
            br.buf = 0;
            br.bits = 0;
            if ((size_t)(br.end - br.in) < len || (size_t)(out_end - out) < len) { refused = true; return; }
            memcpy(out, br.in, len);
            br.in += len;
            out += len;
            return;
        }
        t = &fixed_tables();
        if (type == 2) {
            if (!read_dynamic_tables(br, dyn)) { refused = true; return; }
            t = &dyn;
        }
        in_block = true;
    }

    // one symbol with every bound checked (the ends of the input and of the output)
    void checked_symbol()
    {
        br.refill();
        uint32_t e = t->ll[br.peek(LL_BITS)];
        if ((e & 0x70u) == (SUBTABLE << 4)) {
            br.drop(LL_BITS);
            e = t->ll[(e >> 16) + br.peek((int)((e >> 8) & 0xFFu))];
        }
        br.drop((int)(e & 15u));
        const uint32_t kind = (e >> 4) & 7u;
        if (kind == LITERAL) {
            if (out == out_end) { refused = true; return; }
            *out++ = (uint8_t)(e >> 16);
            return;
        }
        if (kind == END_OF_BLOCK) {
            if (br.bits < 0) refused = true;
            end_block();
            return;
        }
        if (kind != BASE) { refused = true; return; }
        const uint32_t len = (e >> 16) + br.take((int)((e >> 8) & 0xFFu));
        br.refill();
        uint32_t d = t->d[br.peek(D_BITS)];
        if ((d & 0x70u) == (SUBTABLE << 4)) {
            br.drop(D_BITS);
            d = t->d[(d >> 16) + br.peek((int)((d >> 8) & 0xFFu))];
        }
        br.drop((int)(d & 15u));
        if ((d & 0x70u) != (BASE << 4)) { refused = true; return; }
        const uint32_t dist = (d >> 16) + br.take((int)((d >> 8) & 0xFFu));
        if (br.bits < 0 || dist > (size_t)(out - out0) || len > (size_t)(out_end - out)) { refused = true; return; }
        const uint8_t* src = out - dist;
        for (uint8_t* const stop = out + len; out < stop;) *out++ = *src++;
    }

    // until the stream stands in a Huffman block with room for the fast loop, or is over
    void prepare()
    {
        while (live()) {
            if (!in_block) {
                if (last) {
                    if ((out == out_end || !exact) && br.bits >= 0) done = true;
                    else refused = true;
                    return;
                }
                open_block();
                continue;
            }
            if (fast_ok()) return;
            checked_symbol();
        }
    }

    // the rest of the stream, alone
    PMB_INLINE bool finish()
    {
        for (;;) {
            prepare();
            if (!live()) return done;
            Hot h = hot();
            Round r = MORE;
            // (locals: the byte stores of the loop could alias anything reachable from `this`)
            const uint8_t* const in_lim = br.end - FAST_IN;
            uint8_t* const out_lim = out_end - FAST_OUT;
            const Tables* const tables = t;
            const uint8_t* const base = out0;
            while (r == MORE && h.in <= in_lim && h.out <= out_lim) r = fast_round(h, tables, base);
            keep(h);
            if (r == REFUSED) return false;
            if (r == BLOCK_END) end_block();
        }
    }
};

// Inflate in[0, in_len) into exactly out[0, out_len).  false: not decoded (use zlib).
PMB_INLINE bool inflate(const uint8_t* in, size_t in_len, uint8_t* out, size_t out_len)
{
    Decoder d;
    d.init(in, in_len, out, out_len);
    return d.finish();
}

// Inflate in[0, in_len) into out[0, capacity): a stream whose inflated size is not known up
// front (BigWig data blocks).  *produced = its size.  false: not decoded, or it did not fit.
PMB_INLINE bool inflate_upto(const uint8_t* in, size_t in_len, uint8_t* out, size_t capacity, size_t* produced)
{
    Decoder d;
    d.init(in, in_len, out, capacity);
    d.exact = false;
    if (!d.finish()) return false;
    *produced = (size_t)(d.out - d.out0);
    return true;
}

// Two independent streams at once (two BGZF blocks): their fast loops take turns, round by
// round, so that the two dependency chains overlap.  ok[i]: stream i decoded.
PMB_INLINE void inflate2(const uint8_t* const in[2], const size_t in_len[2], uint8_t* const out[2], const size_t out_len[2],
                     bool ok[2])
{
    Decoder a, b;
    a.init(in[0], in_len[0], out[0], out_len[0]);
    b.init(in[1], in_len[1], out[1], out_len[1]);
    for (;;) {
        a.prepare();
        b.prepare();
        if (!a.live() || !b.live()) break;
        Hot ha = a.hot(), hb = b.hot();
        Round ra = MORE, rb = MORE;
        const uint8_t* const a_in_lim = a.br.end - FAST_IN;
        const uint8_t* const b_in_lim = b.br.end - FAST_IN;
        uint8_t* const a_out_lim = a.out_end - FAST_OUT;
        uint8_t* const b_out_lim = b.out_end - FAST_OUT;
        const Tables* const ta = a.t;
        const Tables* const tb = b.t;
        const uint8_t* const a_base = a.out0;
        const uint8_t* const b_base = b.out0;
        while (ha.in <= a_in_lim && ha.out <= a_out_lim && hb.in <= b_in_lim && hb.out <= b_out_lim) {
            ra = fast_round(ha, ta, a_base);
            rb = fast_round(hb, tb, b_base);
            if (ra | rb) break;
        }
        a.keep(ha);
        b.keep(hb);
        if (ra == REFUSED) a.refused = true;
        if (rb == REFUSED) b.refused = true;
        if (ra == BLOCK_END) a.end_block();
        if (rb == BLOCK_END) b.end_block();
    }
    ok[0] = a.refused ? false : a.finish();
    ok[1] = b.refused ? false : b.finish();
}

#undef PMB_INLINE

}  // namespace pmb_inflate
#endif
