// pmb_io.cpp -- libpymasc_io.so (include/pymasc_io.h): BAM (BGZF) and BigWig decoding on the
// host into packed arrays.  zlib + std::thread; no CUDA here.
//
// BAM: the file is memory-mapped, BGZF block headers are walked sequentially (18 bytes per
// <= 64 KiB block), cache-sized batches of blocks are inflated in parallel straight into their
// final offsets of one buffer (ISIZE in each block's footer gives the offsets up front), record
// boundaries are found by one hop over the block_size fields, and the batch is filtered in
// parallel shares whose outputs stay as the handle's storage (no merge copy); a record cut by
// the batch boundary is carried into the next batch.  The reads of every reference are then
// reduced by the calculator's feed rules, references in parallel.
// BigWig: bbi header, chromosome B+ tree and the R-tree index are read from the mapping; a
// chromosome's data blocks are found through the R-tree, inflated and decoded (bedGraph,
// varStep and fixedStep sections).
#include <fcntl.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#include <zlib.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <map>
#include <memory>
#include <string>
#include <thread>
#include <vector>

#include "../../include/pymasc_io.h"
#include "pmb_inflate.h"

static thread_local std::string g_err;

extern "C" const char* pmio_last_error(void) { return g_err.c_str(); }

static int fail(int code, const std::string& msg)
{
    g_err = msg;
    return code;
}

namespace {

struct Mapping {
    const uint8_t* p = nullptr;
    size_t n = 0;
    int fd = -1;
    int open(const char* path)
    {
        fd = ::open(path, O_RDONLY);
        if (fd < 0) return fail(PMIO_ERR_IO, std::string("cannot open ") + path);
        struct stat st;
        if (fstat(fd, &st) != 0) return fail(PMIO_ERR_IO, std::string("cannot stat ") + path);
        n = (size_t)st.st_size;
        if (n == 0) return fail(PMIO_ERR_FORMAT, std::string(path) + " is empty");
        void* m = mmap(nullptr, n, PROT_READ, MAP_PRIVATE, fd, 0);
        if (m == MAP_FAILED) return fail(PMIO_ERR_IO, std::string("cannot map ") + path);
        p = (const uint8_t*)m;
        madvise(m, n, MADV_SEQUENTIAL);
        return PMIO_OK;
    }
    ~Mapping()
    {
        if (p) munmap((void*)p, n);
        if (fd >= 0) ::close(fd);
    }
};

inline uint16_t rd16(const uint8_t* p) { uint16_t v; memcpy(&v, p, 2); return v; }
inline uint32_t rd32(const uint8_t* p) { uint32_t v; memcpy(&v, p, 4); return v; }
inline int32_t rdi32(const uint8_t* p) { int32_t v; memcpy(&v, p, 4); return v; }
inline uint64_t rd64(const uint8_t* p) { uint64_t v; memcpy(&v, p, 8); return v; }
inline float rdf32(const uint8_t* p) { float v; memcpy(&v, p, 4); return v; }

bool zlib_inflate_raw(const uint8_t* src, size_t csize, uint8_t* dst, size_t isize)
{
    z_stream zs;
    memset(&zs, 0, sizeof zs);
    if (inflateInit2(&zs, -15) != Z_OK) return false;
    zs.next_in = (Bytef*)src;
    zs.avail_in = (uInt)csize;
    zs.next_out = dst;
    zs.avail_out = (uInt)isize;
    int rc = inflate(&zs, Z_FINISH);
    const bool ok = rc == Z_STREAM_END && zs.total_out == isize;
    inflateEnd(&zs);
    return ok;
}

// raw-deflate payload -> dst (exactly `isize` bytes); crc checked.  The fast decoder
// (pmb_inflate.h) goes first; whatever it refuses or gets wrong (size or CRC) is done by zlib.
bool inflate_block(const uint8_t* src, size_t csize, uint8_t* dst, size_t isize, uint32_t crc)
{
    if (isize == 0) return true;
    static const bool zlib_only = getenv("PMIO_ZLIB_ONLY") != nullptr;      // A/B aid
    auto crc_ok = [&]() { return (uint32_t)crc32(crc32(0L, Z_NULL, 0), dst, (uInt)isize) == crc; };
    if (!zlib_only && pmb_inflate::inflate(src, csize, dst, isize) && crc_ok()) return true;
    return zlib_inflate_raw(src, csize, dst, isize) && crc_ok();
}

struct Block { size_t data_off; uint32_t csize, isize, crc; size_t out_off; };

}  // namespace

// Raw-DEFLATE decoding of one buffer by either decoder (test hook: the fast decoder is checked
// against zlib on arbitrary streams, tests/test_io_native.py).
extern "C" int pmio_inflate_raw(const uint8_t* in, uint64_t in_len, uint8_t* out, uint64_t out_len, int use_zlib)
{
    if (!in || (!out && out_len)) return fail(PMIO_ERR_ARG, "null argument");
    const bool ok = use_zlib ? zlib_inflate_raw(in, (size_t)in_len, out, (size_t)out_len)
                             : pmb_inflate::inflate(in, (size_t)in_len, out, (size_t)out_len);
    return ok ? PMIO_OK : fail(PMIO_ERR_FORMAT, "stream does not inflate to the given size");
}

// ====================================================================== BAM
// One thread's share of a batch: a run of complete records, filtered into arrays that stay
// alive as the storage of the handle (no merge pass: a reference's reads are the concatenation
// of its segments, in file order).  Accepted reads form runs of one reference (a sorted file has
// one or two per share).
struct pmio_chunk {
    struct Run { int32_t ref; size_t begin; uint32_t first_pos, last_pos; };
    std::vector<uint32_t> pos, qlen;
    std::vector<uint8_t> rev;
    std::vector<Run> runs;
};

struct pmio_bam {
    std::vector<std::string> names;
    std::vector<uint64_t> lens;
    std::vector<std::unique_ptr<pmio_chunk>> store;
    struct Seg { const pmio_chunk* c; size_t begin, end; };
    std::vector<std::vector<Seg>> segs;         // per reference, file order
    std::vector<uint64_t> counts;               // accepted reads per reference
    std::vector<uint32_t> first_pos, last_pos;
    // per-read arrays, concatenated on first request (pmio_bam_reads)
    std::vector<std::vector<uint32_t>> pos, qlen;
    std::vector<std::vector<uint8_t>> rev;
    std::vector<uint8_t> materialised;
    std::vector<uint32_t> hist_len;
    std::vector<uint64_t> hist_cnt;
    pmio_bam_stats st{};
    // per reference, the reads reduced by the calculator's feed rules (mscc.pyx:370-418)
    struct Packed {
        std::vector<uint32_t> fwd, rev;
        uint64_t fwd_len_sum = 0, rev_len_sum = 0;
    };
    std::vector<Packed> packed;
};

namespace {

// Header (magic, text, reference list): sequential, resumable across batches.
struct BamHeaderParser {
    pmio_bam* b;
    int stage = 0;              // 0 magic + text, 1 references, 2 done (records follow)
    uint32_t refs_left = 0;

    // consumes what [p, p+n) holds of the header; returns bytes consumed, or -1
    long long feed(const uint8_t* p, size_t n)
    {
        size_t o = 0;
        if (stage == 0) {
            if (n < 12) return 0;
            if (memcmp(p, "BAM\1", 4) != 0) { g_err = "not a BAM file (bad magic)"; return -1; }
            const int32_t l_text = rdi32(p + 4);
            if (l_text < 0) { g_err = "corrupt BAM header"; return -1; }
            if (n < 12 + (size_t)l_text) return 0;
            refs_left = rd32(p + 8 + l_text);
            o = 12 + (size_t)l_text;
            b->names.reserve(std::min<uint32_t>(refs_left, 1u << 20));
            stage = 1;
        }
        while (stage == 1) {
            if (refs_left == 0) {
                const size_t nr = b->names.size();
                b->segs.resize(nr); b->counts.assign(nr, 0);
                b->first_pos.assign(nr, 0); b->last_pos.assign(nr, 0);
                b->pos.resize(nr); b->qlen.resize(nr); b->rev.resize(nr);
                b->materialised.assign(nr, 0);
                stage = 2;
                break;
            }
            if (n - o < 4) return (long long)o;
            const int32_t l_name = rdi32(p + o);
            if (l_name <= 0) { g_err = "corrupt BAM reference list"; return -1; }
            if (n - o < 8 + (size_t)l_name) return (long long)o;
            b->names.emplace_back((const char*)p + o + 4, strnlen((const char*)p + o + 4, l_name));
            b->lens.push_back(rd32(p + o + 4 + l_name));
            o += 8 + (size_t)l_name;
            refs_left--;
        }
        return (long long)o;
    }
};

// Counters of one share (scratch, reused from batch to batch).
struct ShareStats {
    pmio_bam_stats st{};
    std::vector<uint64_t> hist_small = std::vector<uint64_t>(1024, 0);
    std::map<uint32_t, uint64_t> hist;
    bool sorted = true;         // within every run
    bool incomplete = false;    // stopped at a record that runs past the end of the batch
    const char* err = nullptr;
    void reset()
    {
        incomplete = false;
        st = pmio_bam_stats{};
        std::fill(hist_small.begin(), hist_small.end(), 0);
        hist.clear();
        sorted = true;
        err = nullptr;
    }
};

// Filters the records that START in [start, limit) of the batch buffer p[0, end); returns where
// it stopped: the start of the first record at or beyond `limit`, or of a record that does not
// end inside the batch (ss.incomplete).  `start` may be a guess (a BGZF block boundary): every
// read is bounds-checked, and what a wrong guess produces (ss.err, or a stop that the caller's
// validation rejects) is thrown away.
size_t parse_share(const uint8_t* p, size_t start, size_t limit, size_t end, uint32_t n_refs, int mapq,
                   pmio_chunk& out, ShareStats& ss)
{
    pmio_bam_stats& st = ss.st;
    const size_t guess = (limit - start) / 64 + 16;       // virtual reservation; untouched pages cost nothing
    out.pos.reserve(guess); out.qlen.reserve(guess); out.rev.reserve(guess);
    size_t o = start;
    while (o < limit) {
        if (end - o < 4) { ss.incomplete = true; return o; }
        const int32_t bs = rdi32(p + o);
        if (bs < 32) { ss.err = "corrupt BAM record (block_size < 32)"; return o; }
        if (end - o < 4 + (size_t)bs) { ss.incomplete = true; return o; }
        const uint8_t* r = p + o + 4;
        const size_t rec = o;
        o += 4 + (size_t)bs;
        st.n_records++;
        const int32_t ref = rdi32(r);
        if (ref < 0) continue;                                   // reference_name is None
        if ((uint32_t)ref >= n_refs) { ss.err = "BAM record names an unknown reference"; return rec; }
        const int32_t pos0 = rdi32(r + 4);
        const uint32_t l_name = r[8], mq = r[9];
        const uint32_t n_cig = rd16(r + 12), flag = rd16(r + 14);
        st.n_with_ref++;
        if (flag & 0x1) { st.n_paired++; if (flag & 0x80) st.n_read2++; }
        if (flag & 0x4) st.n_unmapped++;
        if (32 + l_name + 4ull * n_cig > (uint32_t)bs) { ss.err = "corrupt BAM record (CIGAR overruns)"; return rec; }
        // infer_query_length(): M, I, S, =, X; None (skip) without a CIGAR
        uint32_t ql = 0;
        const uint8_t* cg = r + 32 + l_name;
        for (uint32_t i = 0; i < n_cig; i++) {
            const uint32_t c = rd32(cg + 4 * i), op = c & 15;
            if (op == 0 || op == 1 || op == 4 || op == 7 || op == 8) ql += c >> 4;
        }
        const bool good = !(flag & 0x4) && !(flag & 0x400) && (int)mq >= mapq;
        // estimate_readlen's counter (readlen.pyx:150-156): read2 is NOT excluded there
        if (good) {
            if (ql < ss.hist_small.size()) ss.hist_small[ql]++;
            else ss.hist[ql]++;
        }
        if ((flag & 0x80) || !good || ql == 0) continue;        // handler/read.py:62-90, :132-143
        const uint32_t p1 = (uint32_t)(pos0 + 1);
        if (out.runs.empty() || out.runs.back().ref != ref) out.runs.push_back({ref, out.pos.size(), p1, p1});
        else if (p1 < out.runs.back().last_pos) ss.sorted = false;
        out.runs.back().last_pos = p1;
        out.pos.push_back(p1);
        out.qlen.push_back(ql);
        out.rev.push_back((flag & 0x10) ? 1 : 0);
        st.n_accepted++;
    }
    return o;
}

// Shares of a batch, in file order: their runs become segments of the per-reference lists;
// coordinate order across shares and references is checked here (mscc.pyx:351-366: no reference
// revisited, positions non-decreasing).
struct BamMerger {
    pmio_bam* b;
    std::map<uint32_t, uint64_t> hist;
    std::vector<uint64_t> hist_small = std::vector<uint64_t>(1024, 0);
    int32_t cur_ref = -1;
    uint32_t last_pos = 0;
    std::vector<uint8_t> seen;
    bool sorted = true;

    void take(const pmio_chunk& c, const ShareStats& ss)
    {
        pmio_bam_stats& st = b->st;
        st.n_records += ss.st.n_records; st.n_with_ref += ss.st.n_with_ref; st.n_unmapped += ss.st.n_unmapped;
        st.n_paired += ss.st.n_paired; st.n_read2 += ss.st.n_read2; st.n_accepted += ss.st.n_accepted;
        for (size_t l = 0; l < hist_small.size(); l++) hist_small[l] += ss.hist_small[l];
        for (auto& kv : ss.hist) hist[kv.first] += kv.second;
        if (!ss.sorted) sorted = false;
        if (seen.size() < b->names.size()) seen.resize(b->names.size(), 0);
        for (size_t k = 0; k < c.runs.size(); k++) {
            const pmio_chunk::Run& r = c.runs[k];
            const size_t end = k + 1 < c.runs.size() ? c.runs[k + 1].begin : c.pos.size();
            if (r.ref != cur_ref) {
                if (seen[r.ref]) sorted = false;
                seen[r.ref] = 1;
                cur_ref = r.ref;
                last_pos = 0;
            }
            if (r.first_pos < last_pos) sorted = false;
            last_pos = r.last_pos;
            if (b->counts[r.ref] == 0) b->first_pos[r.ref] = r.first_pos;
            b->last_pos[r.ref] = r.last_pos;
            b->counts[r.ref] += end - r.begin;
            b->segs[r.ref].push_back({&c, r.begin, end});
        }
    }
};

// What feed_forward_read / feed_reverse_read leave behind for one chromosome of sorted reads:
// forward 5' positions with consecutive repeats dropped (mscc.pyx:388-393: a read at
// _last_forward_pos is a duplicate) and the query lengths of the kept reads summed; reverse 5'
// ends pos + qlen - 1, each bit counted once with the query length of the FIRST read that set
// it (mscc.pyx:416-418).  Both lists come out sorted and unique.
void pack_reference(const std::vector<pmio_bam::Seg>& segs, uint64_t n, pmio_bam::Packed& out)
{
    out.fwd.reserve(n / 2 + n / 8 + 16);
    std::vector<uint32_t> ends, lens;
    ends.reserve(n / 2 + n / 8 + 16);
    lens.reserve(n / 2 + n / 8 + 16);
    uint32_t last_f = 0;
    bool ends_sorted = true;
    for (const pmio_bam::Seg& s : segs) {
        const uint32_t* pos = s.c->pos.data();
        const uint32_t* qlen = s.c->qlen.data();
        const uint8_t* rev = s.c->rev.data();
        for (size_t i = s.begin; i < s.end; i++) {
            if (!rev[i]) {
                if (pos[i] == last_f) continue;
                last_f = pos[i];
                out.fwd.push_back(pos[i]);
                out.fwd_len_sum += qlen[i];
            } else {
                const uint32_t e = pos[i] + qlen[i] - 1;
                if (!ends.empty() && e < ends.back()) ends_sorted = false;
                ends.push_back(e);
                lens.push_back(qlen[i]);
            }
        }
    }
    if (ends_sorted) {
        out.rev.reserve(ends.size());
        for (size_t i = 0; i < ends.size(); i++)
            if (i == 0 || ends[i] != ends[i - 1]) { out.rev.push_back(ends[i]); out.rev_len_sum += lens[i]; }
        return;
    }
    std::vector<uint32_t> idx(ends.size());
    for (size_t i = 0; i < idx.size(); i++) idx[i] = (uint32_t)i;
    std::stable_sort(idx.begin(), idx.end(), [&](uint32_t a, uint32_t b) { return ends[a] < ends[b]; });
    for (size_t k = 0; k < idx.size(); k++)
        if (k == 0 || ends[idx[k]] != ends[idx[k - 1]]) { out.rev.push_back(ends[idx[k]]); out.rev_len_sum += lens[idx[k]]; }
}

}  // namespace

static int bam_load_impl(const char* path, int mapq_criteria, int threads, pmio_bam** out);

// Nothing may propagate out of the C ABI: allocation failures and anything unexpected become codes.
template <class F>
static int guarded(F&& body)
{
    try {
        return body();
    } catch (const std::bad_alloc&) {
        return fail(PMIO_ERR_NOMEM, "out of memory");
    } catch (const std::exception& e) {
        return fail(PMIO_ERR_FORMAT, std::string("unexpected error: ") + e.what());
    } catch (...) {
        return fail(PMIO_ERR_FORMAT, "unexpected error");
    }
}

extern "C" int pmio_bam_load(const char* path, int mapq_criteria, int threads, pmio_bam** out)
{
    if (!path || !out) return fail(PMIO_ERR_ARG, "null argument");
    *out = nullptr;
    return guarded([&]() { return bam_load_impl(path, mapq_criteria, threads, out); });
}

static int bam_load_impl(const char* path, int mapq_criteria, int threads, pmio_bam** out)
{
    Mapping m;
    int rc = m.open(path);
    if (rc != PMIO_OK) return rc;
    if (threads <= 0) threads = (int)std::max(1u, std::thread::hardware_concurrency());
    threads = std::min(threads, 64);

    std::unique_ptr<pmio_bam> owner(new pmio_bam());      // freed on every early return and on exceptions
    pmio_bam* b = owner.get();
    BamHeaderParser header{b};
    BamMerger merger{b};
    std::vector<ShareStats> share_stats((size_t)threads);
    std::vector<size_t> cuts, stops;
    // PMIO_TIMING=1: seconds per stage on stderr (debug aid)
    static const bool timing = getenv("PMIO_TIMING") != nullptr;
    auto now = []() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    double t_inflate = 0, t_scan = 0, t_parse = 0, t_pack = 0, t_mark = now();
    // Batch size: large enough that the two thread-pool start-ups per batch are noise (measured:
    // 256-block batches on 8 threads cost 1.7x the inflate time of 4096-block ones), small
    // enough to bound the buffer (<= 256 MiB inflated).
    size_t BATCH_BLOCKS = std::min<size_t>(4096, 512 * (size_t)threads);
    if (const char* e = getenv("PMIO_BATCH_BLOCKS")) BATCH_BLOCKS = std::max<size_t>(1, (size_t)atoll(e));
    std::vector<uint8_t> buf;
    std::vector<Block> blocks;
    size_t carry = 0, off = 0;
    bool bad = false;
    while (off < m.n && !bad) {
        // ---- collect a batch of block headers
        blocks.clear();
        size_t out_off = carry;
        while (off < m.n && blocks.size() < BATCH_BLOCKS) {
            if (m.n - off < 18 || m.p[off] != 31 || m.p[off + 1] != 139 || m.p[off + 2] != 8 || !(m.p[off + 3] & 4)) {
                g_err = blocks.empty() && off == 0 ? "not a BGZF (BAM) file" : "corrupt BGZF block header";
                bad = true;
                break;
            }
            const uint32_t xlen = rd16(m.p + off + 10);
            if (m.n - off < 12 + (size_t)xlen) { g_err = "truncated BGZF block"; bad = true; break; }
            uint32_t bsize = 0;
            for (uint32_t x = 0; x + 4 <= xlen;) {       // find the BC subfield
                const uint8_t* e = m.p + off + 12 + x;
                const uint32_t slen = rd16(e + 2);
                if (e[0] == 'B' && e[1] == 'C' && slen == 2 && x + 6 <= xlen) { bsize = rd16(e + 4) + 1u; break; }
                x += 4 + slen;
            }
            if (bsize < 12 + xlen + 8 || m.n - off < bsize) { g_err = "truncated BGZF block"; bad = true; break; }
            Block k;
            k.data_off = off + 12 + xlen;
            k.csize = bsize - 12 - xlen - 8;
            k.crc = rd32(m.p + off + bsize - 8);
            k.isize = rd32(m.p + off + bsize - 4);
            if (k.isize > 65536) { g_err = "corrupt BGZF block (ISIZE > 64 KiB)"; bad = true; break; }
            k.out_off = out_off;
            out_off += k.isize;
            if (k.isize) blocks.push_back(k);
            off += bsize;
        }
        if (bad) break;
        if (buf.size() < out_off) buf.resize(out_off);
        // ---- inflate the batch in parallel
        std::atomic<size_t> next{0};
        std::atomic<bool> ok{true};
        auto work = [&]() {
            for (;;) {
                const size_t i = next.fetch_add(1);
                if (i >= blocks.size() || !ok.load()) return;
                const Block& k = blocks[i];
                if (!inflate_block(m.p + k.data_off, k.csize, buf.data() + k.out_off, k.isize, k.crc)) ok.store(false);
            }
        };
        const int nt = (int)std::min<size_t>((size_t)threads, std::max<size_t>(1, blocks.size() / 4));
        std::vector<std::thread> pool;
        for (int t = 1; t < nt; t++) pool.emplace_back(work);
        work();
        for (auto& t : pool) t.join();
        if (!ok.load()) { g_err = "BGZF block does not inflate (corrupt data or CRC mismatch)"; bad = true; break; }
        { const double t = now(); t_inflate += t - t_mark; t_mark = t; }
        // ---- header first (sequential), then the records.  Shares start at BGZF block boundaries
        //      -- htslib never lets a record straddle two blocks unless it is longer than one, so a
        //      block boundary is almost always a record boundary -- and are filtered in parallel ON
        //      THAT GUESS; afterwards the guesses are checked inductively in file order (share 0
        //      starts at a known boundary; share i+1's start is right iff it is where share i
        //      stopped) and a share whose start was wrong is filtered again from the true boundary.
        //      No sequential pass over the records remains in the common case.  The record cut by
        //      the end of the batch is carried into the next one.
        size_t used = 0;
        if (header.stage != 2) {
            const long long h = header.feed(buf.data(), out_off);
            if (h < 0) { bad = true; break; }
            used = (size_t)h;
        }
        if (header.stage == 2) {
            const uint8_t* p = buf.data();
            const size_t share = std::max<size_t>((out_off - used) / (size_t)threads + 1, 1 << 16);
            cuts.clear();
            cuts.push_back(used);
            for (const Block& k : blocks)
                if (k.out_off > cuts.back() && k.out_off - cuts.back() >= share) cuts.push_back(k.out_off);
            cuts.push_back(out_off);
            const size_t n_shares = cuts.size() - 1;
            const uint32_t n_refs = (uint32_t)b->names.size();
            if (share_stats.size() < n_shares) share_stats.resize(n_shares);
            stops.assign(n_shares, 0);
            const size_t store0 = b->store.size();
            for (size_t i = 0; i < n_shares; i++) b->store.emplace_back(new pmio_chunk());
            std::atomic<size_t> next_share{0};
            std::atomic<bool> oom{false};
            auto parse_work = [&]() {
                for (size_t i; (i = next_share.fetch_add(1)) < n_shares;) {
                    share_stats[i].reset();
                    try {       // an exception must not leave a worker thread
                        stops[i] = parse_share(p, cuts[i], cuts[i + 1], out_off, n_refs, mapq_criteria,
                                               *b->store[store0 + i], share_stats[i]);
                    } catch (...) {
                        oom.store(true);
                    }
                }
            };
            std::vector<std::thread> ppool;
            for (size_t t = 1; t < std::min<size_t>((size_t)threads, n_shares); t++) ppool.emplace_back(parse_work);
            parse_work();
            for (auto& t : ppool) t.join();
            if (oom.load()) throw std::bad_alloc();
            { const double t = now(); t_parse += t - t_mark; t_mark = t; }
            size_t at = used;                       // a true record boundary
            for (size_t i = 0; i < n_shares; i++) {
                pmio_chunk& c = *b->store[store0 + i];
                if (at >= cuts[i + 1] && cuts[i + 1] != cuts[i]) { c = pmio_chunk(); continue; }   // inside a long record
                if (at != cuts[i]) {                // wrong guess: again, from the true boundary
                    c = pmio_chunk();
                    share_stats[i].reset();
                    stops[i] = parse_share(p, at, cuts[i + 1], out_off, n_refs, mapq_criteria, c, share_stats[i]);
                    b->st.reparsed_shares++;
                }
                if (share_stats[i].err) { g_err = share_stats[i].err; bad = true; break; }
                merger.take(c, share_stats[i]);
                at = stops[i];
                if (share_stats[i].incomplete) {
                    for (size_t k = i + 1; k < n_shares; k++) *b->store[store0 + k] = pmio_chunk();
                    break;
                }
            }
            if (bad) break;
            used = at;
            { const double t = now(); t_scan += t - t_mark; t_mark = t; }
        }
        carry = out_off - used;
        if (carry && used) memmove(buf.data(), buf.data() + used, carry);
    }
    if (!bad && header.stage != 2) { g_err = "truncated BAM header"; bad = true; }
    if (!bad && carry) { g_err = "truncated BAM file (incomplete record at the end)"; bad = true; }
    if (bad) return PMIO_ERR_FORMAT;
    b->st.sorted = merger.sorted ? 1 : 0;
    for (size_t l = 0; l < merger.hist_small.size(); l++)
        if (merger.hist_small[l]) merger.hist[(uint32_t)l] += merger.hist_small[l];
    for (auto& kv : merger.hist) { b->hist_len.push_back(kv.first); b->hist_cnt.push_back(kv.second); }
    // packed form per reference, references spread over the threads (longest first)
    if (b->st.sorted) {
        b->packed.resize(b->names.size());
        std::vector<size_t> order;
        for (size_t i = 0; i < b->names.size(); i++)
            if (b->counts[i]) order.push_back(i);
        std::sort(order.begin(), order.end(), [&](size_t x, size_t y) { return b->counts[x] > b->counts[y]; });
        std::atomic<size_t> next{0};
        std::atomic<bool> oom{false};
        auto work = [&]() {
            for (size_t k; (k = next.fetch_add(1)) < order.size();) {
                try {
                    pack_reference(b->segs[order[k]], b->counts[order[k]], b->packed[order[k]]);
                } catch (...) {
                    oom.store(true);
                }
            }
        };
        std::vector<std::thread> pool;
        for (size_t t = 1; t < std::min<size_t>((size_t)threads, order.size()); t++) pool.emplace_back(work);
        work();
        for (auto& t : pool) t.join();
        if (oom.load()) throw std::bad_alloc();
    }
    t_pack = now() - t_mark;
    if (timing)
        fprintf(stderr, "[pmio bam] %d threads, %zu-block batches: inflate %.3f s, filter %.3f s, validate + carry %.3f s "
                        "(%u shares filtered again), pack %.3f s\n", threads, BATCH_BLOCKS, t_inflate, t_parse, t_scan,
                b->st.reparsed_shares, t_pack);
    *out = owner.release();
    return PMIO_OK;
}

extern "C" void pmio_bam_free(pmio_bam* b) { delete b; }

extern "C" int pmio_bam_stats_get(const pmio_bam* b, pmio_bam_stats* s)
{
    if (!b || !s) return fail(PMIO_ERR_ARG, "null argument");
    *s = b->st;
    return PMIO_OK;
}

extern "C" uint32_t pmio_bam_n_refs(const pmio_bam* b) { return b ? (uint32_t)b->names.size() : 0; }

extern "C" const char* pmio_bam_ref_name(const pmio_bam* b, uint32_t i)
{
    return b && i < b->names.size() ? b->names[i].c_str() : nullptr;
}

extern "C" uint64_t pmio_bam_ref_len(const pmio_bam* b, uint32_t i)
{
    return b && i < b->lens.size() ? b->lens[i] : 0;
}

extern "C" int pmio_bam_reads(pmio_bam* b, uint32_t i, const uint32_t** pos, const uint32_t** qlen,
                              const uint8_t** is_reverse, uint64_t* n)
{
    if (!b || i >= b->names.size()) return fail(PMIO_ERR_ARG, "reference index out of range");
    if ((pos || qlen || is_reverse) && !b->materialised[i]) {      // concatenate the segments once
      const int rc = guarded([&]() {
        b->pos[i].reserve(b->counts[i]); b->qlen[i].reserve(b->counts[i]); b->rev[i].reserve(b->counts[i]);
        for (const pmio_bam::Seg& s : b->segs[i]) {
            b->pos[i].insert(b->pos[i].end(), s.c->pos.begin() + s.begin, s.c->pos.begin() + s.end);
            b->qlen[i].insert(b->qlen[i].end(), s.c->qlen.begin() + s.begin, s.c->qlen.begin() + s.end);
            b->rev[i].insert(b->rev[i].end(), s.c->rev.begin() + s.begin, s.c->rev.begin() + s.end);
        }
        b->materialised[i] = 1;
        return (int)PMIO_OK;
      });
      if (rc != PMIO_OK) return rc;
    }
    if (pos) *pos = b->pos[i].data();
    if (qlen) *qlen = b->qlen[i].data();
    if (is_reverse) *is_reverse = b->rev[i].data();
    if (n) *n = b->counts[i];
    return PMIO_OK;
}

extern "C" int pmio_bam_packed(const pmio_bam* b, uint32_t i, const uint32_t** fwd, uint64_t* n_fwd,
                               uint64_t* fwd_len_sum, const uint32_t** rev, uint64_t* n_rev,
                               uint64_t* rev_len_sum, uint32_t* first_pos, uint32_t* last_pos)
{
    if (!b || i >= b->names.size()) return fail(PMIO_ERR_ARG, "reference index out of range");
    if (!b->st.sorted) return fail(PMIO_ERR_ARG, "reads are not coordinate sorted");
    const pmio_bam::Packed& k = b->packed[i];
    if (fwd) *fwd = k.fwd.data();
    if (n_fwd) *n_fwd = k.fwd.size();
    if (fwd_len_sum) *fwd_len_sum = k.fwd_len_sum;
    if (rev) *rev = k.rev.data();
    if (n_rev) *n_rev = k.rev.size();
    if (rev_len_sum) *rev_len_sum = k.rev_len_sum;
    if (first_pos) *first_pos = b->first_pos[i];
    if (last_pos) *last_pos = b->last_pos[i];
    return PMIO_OK;
}

extern "C" int pmio_bam_readlen_hist(const pmio_bam* b, const uint32_t** lengths, const uint64_t** counts,
                                     uint32_t* n)
{
    if (!b) return fail(PMIO_ERR_ARG, "null argument");
    if (lengths) *lengths = b->hist_len.data();
    if (counts) *counts = b->hist_cnt.data();
    if (n) *n = (uint32_t)b->hist_len.size();
    return PMIO_OK;
}

// ====================================================================== BigWig
struct pmio_bw {
    Mapping m;
    std::vector<std::string> names;
    std::vector<uint64_t> lens;
    std::vector<uint32_t> ids;
    uint64_t index_off = 0;
    uint32_t uncompress_buf = 0;
    std::vector<uint32_t> begins, ends;
    std::vector<float> values;
};

namespace {

bool in_file(const pmio_bw* w, uint64_t off, uint64_t len) { return off <= w->m.n && len <= w->m.n - off; }

// chromosome B+ tree: collect (name, id, size) of every leaf item
bool walk_chrom_tree(pmio_bw* w, uint64_t node, uint32_t key_size, int depth)
{
    if (depth > 32 || !in_file(w, node, 4)) return false;
    const uint8_t* p = w->m.p + node;
    const bool leaf = p[0] != 0;
    const uint32_t count = rd16(p + 2);
    const uint64_t item = key_size + 8;
    if (!in_file(w, node + 4, count * item)) return false;
    for (uint32_t i = 0; i < count; i++) {
        const uint8_t* it = p + 4 + i * item;
        if (leaf) {
            w->names.emplace_back((const char*)it, strnlen((const char*)it, key_size));
            w->ids.push_back(rd32(it + key_size));
            w->lens.push_back(rd32(it + key_size + 4));
        } else if (!walk_chrom_tree(w, rd64(it + key_size), key_size, depth + 1)) {
            return false;
        }
    }
    return true;
}

struct Leaf { uint64_t off, size; };

// R-tree: data blocks that may hold items of chromosome `id`, in index order
bool walk_rtree(const pmio_bw* w, uint64_t node, uint32_t id, std::vector<Leaf>& out, int depth)
{
    if (depth > 32 || !in_file(w, node, 4)) return false;
    const uint8_t* p = w->m.p + node;
    const bool leaf = p[0] != 0;
    const uint32_t count = rd16(p + 2);
    const uint64_t item = leaf ? 32 : 24;
    if (!in_file(w, node + 4, count * item)) return false;
    for (uint32_t i = 0; i < count; i++) {
        const uint8_t* it = p + 4 + i * item;
        const uint32_t c0 = rd32(it), c1 = rd32(it + 8);
        if (id < c0 || id > c1) continue;
        if (leaf) out.push_back({rd64(it + 16), rd64(it + 24)});
        else if (!walk_rtree(w, rd64(it + 16), id, out, depth + 1)) return false;
    }
    return true;
}

}  // namespace

static int bw_open_impl(const char* path, pmio_bw** out);
extern "C" int pmio_bw_open(const char* path, pmio_bw** out)
{
    if (!path || !out) return fail(PMIO_ERR_ARG, "null argument");
    *out = nullptr;
    return guarded([&]() { return bw_open_impl(path, out); });
}

static int bw_open_impl(const char* path, pmio_bw** out)
{
    std::unique_ptr<pmio_bw> owner(new pmio_bw());
    pmio_bw* w = owner.get();
    int rc = w->m.open(path);
    if (rc != PMIO_OK) return rc;
    const uint8_t* p = w->m.p;
    if (w->m.n < 64 || rd32(p) != 0x888FFC26u) {
        return fail(PMIO_ERR_FORMAT, std::string(path) + " is not a little-endian BigWig file (bad magic)");
    }
    const uint64_t chrom_tree = rd64(p + 8);
    w->index_off = rd64(p + 24);
    w->uncompress_buf = rd32(p + 52);
    if (!in_file(w, chrom_tree, 32) || rd32(p + chrom_tree) != 0x78CA8C91u ||
        !in_file(w, w->index_off, 48) || rd32(p + w->index_off) != 0x2468ACE0u) {
        return fail(PMIO_ERR_FORMAT, "corrupt BigWig file (chromosome tree or index)");
    }
    const uint32_t key_size = rd32(p + chrom_tree + 8), val_size = rd32(p + chrom_tree + 12);
    if (val_size != 8 || key_size == 0 || key_size > 4096 || !walk_chrom_tree(w, chrom_tree + 32, key_size, 0)) {
        return fail(PMIO_ERR_FORMAT, "corrupt BigWig chromosome tree");
    }
    *out = owner.release();
    return PMIO_OK;
}

extern "C" void pmio_bw_close(pmio_bw* w) { delete w; }
extern "C" uint32_t pmio_bw_n_chroms(const pmio_bw* w) { return w ? (uint32_t)w->names.size() : 0; }
extern "C" const char* pmio_bw_chrom_name(const pmio_bw* w, uint32_t i)
{
    return w && i < w->names.size() ? w->names[i].c_str() : nullptr;
}
extern "C" uint64_t pmio_bw_chrom_len(const pmio_bw* w, uint32_t i)
{
    return w && i < w->lens.size() ? w->lens[i] : 0;
}

static int bw_intervals_impl(pmio_bw* w, const char* chrom, float threshold, const uint32_t** begins,
                             const uint32_t** ends, const float** values, uint64_t* n);
extern "C" int pmio_bw_intervals(pmio_bw* w, const char* chrom, float threshold, const uint32_t** begins,
                                 const uint32_t** ends, const float** values, uint64_t* n)
{
    if (!w || !chrom) return fail(PMIO_ERR_ARG, "null argument");
    return guarded([&]() { return bw_intervals_impl(w, chrom, threshold, begins, ends, values, n); });
}

static int bw_intervals_impl(pmio_bw* w, const char* chrom, float threshold, const uint32_t** begins,
                             const uint32_t** ends, const float** values, uint64_t* n)
{
    size_t ci = 0;
    for (; ci < w->names.size(); ci++)
        if (w->names[ci] == chrom) break;
    if (ci == w->names.size()) return fail(PMIO_ERR_KEY, chrom);
    const uint32_t id = w->ids[ci];
    const uint64_t chrom_len = w->lens[ci];
    std::vector<Leaf> leaves;
    if (!walk_rtree(w, w->index_off + 48, id, leaves, 0)) return fail(PMIO_ERR_FORMAT, "corrupt BigWig index");
    w->begins.clear(); w->ends.clear(); w->values.clear();
    std::vector<uint8_t> buf;
    for (const Leaf& lf : leaves) {
        if (!in_file(w, lf.off, lf.size)) return fail(PMIO_ERR_FORMAT, "BigWig data block outside the file");
        const uint8_t* d = w->m.p + lf.off;
        size_t dn = (size_t)lf.size;
        if (w->uncompress_buf) {
            buf.resize(std::max<size_t>(w->uncompress_buf, 64));
            uLongf got = (uLongf)buf.size();
            int zr = uncompress(buf.data(), &got, d, (uLong)lf.size);
            while (zr == Z_BUF_ERROR && buf.size() < (1u << 28)) {       // header understated the size
                buf.resize(buf.size() * 2);
                got = (uLongf)buf.size();
                zr = uncompress(buf.data(), &got, d, (uLong)lf.size);
            }
            if (zr != Z_OK) return fail(PMIO_ERR_FORMAT, "BigWig data block does not inflate");
            d = buf.data();
            dn = (size_t)got;
        }
        if (dn < 24) return fail(PMIO_ERR_FORMAT, "short BigWig data block");
        if (rd32(d) != id) continue;
        const uint32_t start0 = rd32(d + 4), step = rd32(d + 12), span = rd32(d + 16);
        const uint32_t type = d[20], count = rd16(d + 22);
        const size_t item = type == 1 ? 12 : type == 2 ? 8 : 4;
        if (type < 1 || type > 3 || dn < 24 + count * item) return fail(PMIO_ERR_FORMAT, "corrupt BigWig data block");
        for (uint32_t i = 0; i < count; i++) {
            const uint8_t* it = d + 24 + i * item;
            uint32_t s, e;
            float v;
            if (type == 1) { s = rd32(it); e = rd32(it + 4); v = rdf32(it + 8); }
            else if (type == 2) { s = rd32(it); e = s + span; v = rdf32(it + 4); }
            else { s = start0 + i * step; e = s + span; v = rdf32(it); }
            if (s >= chrom_len || e == 0) continue;                  // intervals(chrom, 0, size): overlap only
            if (threshold > 0 && !(v >= threshold)) continue;        // reader/bigwig.pyx:174-175
            w->begins.push_back(s); w->ends.push_back(e); w->values.push_back(v);
        }
    }
    if (begins) *begins = w->begins.data();
    if (ends) *ends = w->ends.data();
    if (values) *values = w->values.data();
    if (n) *n = w->begins.size();
    return PMIO_OK;
}
