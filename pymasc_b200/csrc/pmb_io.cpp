// pmb_io.cpp -- libpymasc_io.so (include/pymasc_io.h): BAM (BGZF) and BigWig decoding on the
// host into packed arrays.  Own inflate / CRC-32 (pmb_inflate.h, pmb_crc32.h) with zlib behind
// them, std::thread; no CUDA here.
//
// BAM: the file is memory-mapped and the BGZF block headers are walked sequentially (18 bytes per
// <= 64 KiB block; ISIZE in each footer gives every block's offset in the inflated stream up
// front).  A batch of blocks (cache-sized, 64-256 MiB inflated) is cut into shares of about 512 KiB; a worker
// thread inflates one share (pmb_inflate.h, CRC-checked, zlib behind it) and filters its records
// straight away, while they are in its cache, into arrays that stay as the handle's storage (no
// merge copy).  Where a share's first record starts is guessed from the bytes (the block
// boundary in htslib-written files, a few bytes on in htsjdk-written ones) and the guesses are
// validated in file order afterwards; the one record that crosses a share's end is filtered
// then, and a record cut by the end of the batch is carried into the next.  The reads of every
// reference are finally reduced by the calculator's feed rules, references in parallel.
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
#include <condition_variable>
#include <functional>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/pymasc_io.h"
#include "pmb_crc32.h"
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

// The decoder's entry points, compiled twice: for any x86-64 and for BMI2 (SHRX / SHLX / BZHI in
// the bit-buffer arithmetic: +9 % per thread); the loader picks one (ifunc).
#if defined(__x86_64__) && defined(__GNUC__) && !defined(__clang__) && !defined(__SANITIZE_THREAD__) && !defined(__SANITIZE_ADDRESS__)
#define PMB_CLONES __attribute__((target_clones("default", "bmi2")))
#else
#define PMB_CLONES
#endif
PMB_CLONES bool own_inflate(const uint8_t* in, size_t in_len, uint8_t* out, size_t out_len)
{
    return pmb_inflate::inflate(in, in_len, out, out_len);
}
PMB_CLONES bool own_inflate_upto(const uint8_t* in, size_t in_len, uint8_t* out, size_t capacity, size_t* produced)
{
    return pmb_inflate::inflate_upto(in, in_len, out, capacity, produced);
}
PMB_CLONES void own_inflate2(const uint8_t* const in[2], const size_t in_len[2], uint8_t* const out[2],
                             const size_t out_len[2], bool ok[2])
{
    pmb_inflate::inflate2(in, in_len, out, out_len, ok);
}

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
    auto crc_ok = [&]() { return pmb_crc32::compute(dst, isize, !zlib_only) == crc; };
    if (!zlib_only && own_inflate(src, csize, dst, isize) && crc_ok()) return true;
    return zlib_inflate_raw(src, csize, dst, isize) && crc_ok();
}

// Two blocks at once: their decode loops take turns (pmb_inflate::inflate2), each is CRC-checked,
// and one the fast decoder refuses or gets wrong goes through zlib like a single block.
bool inflate_block_pair(const uint8_t* const src[2], const size_t csize[2], uint8_t* const dst[2], const size_t isize[2],
                        const uint32_t crc[2])
{
    static const bool zlib_only = getenv("PMIO_ZLIB_ONLY") != nullptr;
    bool ok[2] = {false, false};
    if (!zlib_only && isize[0] && isize[1]) own_inflate2(src, csize, dst, isize, ok);
    for (int i = 0; i < 2; i++) {
        if (ok[i] && pmb_crc32::compute(dst[i], isize[i], true) == crc[i]) continue;
        if (!inflate_block(src[i], csize[i], dst[i], isize[i], crc[i])) return false;
    }
    return true;
}

struct Block { size_t data_off; uint32_t csize, isize, crc; size_t out_off; };

// Threads that live as long as one pmio_bam_load / pmio_bw_prefetch: a BAM file is decoded in many
// small batches (the buffer stays in cache), and starting threads per batch would cost more than
// the batch.
// run(f, helpers): f() on the caller and on up to `helpers` pool threads; returns when all are back.
class WorkerPool {
    std::vector<std::thread> threads;
    std::mutex mu;
    std::condition_variable cv_go, cv_done;
    std::function<void()> job;
    uint64_t generation = 0;
    size_t want = 0, taken = 0, running = 0;
    bool quit = false;

    void loop()
    {
        uint64_t seen = 0;
        std::unique_lock<std::mutex> lk(mu);
        for (;;) {
            cv_go.wait(lk, [&] { return quit || (generation != seen && taken < want); });
            if (quit) return;
            seen = generation;
            taken++;
            std::function<void()> f = job;
            lk.unlock();
            f();
            lk.lock();
            if (--running == 0) cv_done.notify_one();
        }
    }

public:
    explicit WorkerPool(size_t n)
    {
        for (size_t i = 0; i < n; i++) threads.emplace_back([this] { loop(); });
    }
    ~WorkerPool()
    {
        { std::lock_guard<std::mutex> lk(mu); quit = true; }
        cv_go.notify_all();
        for (auto& t : threads) t.join();
    }
    void run(const std::function<void()>& f, size_t helpers)
    {
        helpers = std::min(helpers, threads.size());
        if (helpers) {
            { std::lock_guard<std::mutex> lk(mu); job = f; want = running = helpers; taken = 0; generation++; }
            cv_go.notify_all();
        }
        f();
        if (helpers) {
            std::unique_lock<std::mutex> lk(mu);
            cv_done.wait(lk, [&] { return running == 0; });
        }
    }
};

}  // namespace

// Raw-DEFLATE decoding of one buffer by either decoder (test hook: the fast decoder is checked
// against zlib on arbitrary streams, tests/test_io_native.py).
extern "C" int pmio_inflate_raw(const uint8_t* in, uint64_t in_len, uint8_t* out, uint64_t out_len, int use_zlib)
{
    if (!in || (!out && out_len)) return fail(PMIO_ERR_ARG, "null argument");
    const bool ok = use_zlib ? zlib_inflate_raw(in, (size_t)in_len, out, (size_t)out_len)
                             : own_inflate(in, (size_t)in_len, out, (size_t)out_len);
    return ok ? PMIO_OK : fail(PMIO_ERR_FORMAT, "stream does not inflate to the given size");
}

extern "C" int pmio_inflate_raw2(const uint8_t* const in[2], const uint64_t in_len[2], uint8_t* const out[2],
                                 const uint64_t out_len[2], int ok[2])
{
    if (!in || !in_len || !out || !out_len || !ok || !in[0] || !in[1] || !out[0] || !out[1]) return fail(PMIO_ERR_ARG, "null argument");
    const size_t il[2] = {(size_t)in_len[0], (size_t)in_len[1]}, ol[2] = {(size_t)out_len[0], (size_t)out_len[1]};
    bool good[2];
    own_inflate2(in, il, out, ol, good);
    ok[0] = good[0];
    ok[1] = good[1];
    return PMIO_OK;
}

extern "C" uint32_t pmio_crc32(const uint8_t* data, uint64_t len, int use_zlib)
{
    return data || !len ? pmb_crc32::compute(data, (size_t)len, !use_zlib) : 0;
}

// ====================================================================== BAM
// One thread's share of a batch: a run of complete records, filtered into arrays that stay
// alive as the storage of the handle (no merge pass: a reference's reads are the concatenation
// of its segments, in file order).  Accepted reads form runs of one reference (a sorted file has
// one or two per share).
// (cache-line aligned, like the other per-thread structs below: neighbours are written by
// different threads once per record)
struct alignas(128) pmio_chunk {
    struct Run { int32_t ref; size_t begin; uint32_t first_pos, last_pos; bool sorted = true; };
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
    std::vector<uint32_t> hist_len;             // in the order the lengths first occur in the file
    std::vector<uint64_t> hist_cnt;
    // coordinate order per reference (the reference checks it only on reads that survive the
    // chromosome filter, handler/calc.py:143-144 -> mscc.pyx:351-366): ref_ok[i] = the accepted
    // reads of reference i have non-decreasing positions in file order; run_refs = the references
    // in the order their reads appear, consecutive repeats merged
    std::vector<uint8_t> ref_ok;
    std::vector<int32_t> run_refs;
    bool run_overflow = false;
    pmio_bam_stats st{};
    // per reference, the reads reduced by the calculator's feed rules (mscc.pyx:370-418)
    struct alignas(128) Packed {
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
                b->ref_ok.assign(nr, 1);
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
struct alignas(128) ShareStats {
    pmio_bam_stats st{};
    std::vector<uint64_t> hist_small = std::vector<uint64_t>(512, 0);
    std::map<uint32_t, uint64_t> hist;
    std::vector<uint32_t> first_seen;   // the share's lengths in the order they first occur (MODE ties, readlen.pyx:81)
    bool sorted = true;         // within every run
    bool incomplete = false;    // stopped at a record that runs past the end of the batch
    const char* err = nullptr;
    void reset()
    {
        incomplete = false;
        st = pmio_bam_stats{};
        std::fill(hist_small.begin(), hist_small.end(), 0);
        hist.clear();
        first_seen.clear();
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
        // estimate_readlen's counter (readlen.pyx:150-156): read2 is NOT excluded there.  A record
        // without a CIGAR has no query length (pysam's infer_query_length() gives None): not counted
        if (good && n_cig) {
            if (ql < ss.hist_small.size()) { if (ss.hist_small[ql]++ == 0) ss.first_seen.push_back(ql); }
            else if (ss.hist[ql]++ == 0) ss.first_seen.push_back(ql);
        }
        if ((flag & 0x80) || !good || ql == 0) continue;        // handler/read.py:62-90, :132-143
        const uint32_t p1 = (uint32_t)(pos0 + 1);
        if (out.runs.empty() || out.runs.back().ref != ref) out.runs.push_back({ref, out.pos.size(), p1, p1});
        else if (p1 < out.runs.back().last_pos) { ss.sorted = false; out.runs.back().sorted = false; }
        out.runs.back().last_pos = p1;
        out.pos.push_back(p1);
        out.qlen.push_back(ql);
        out.rev.push_back((flag & 0x10) ? 1 : 0);
        st.n_accepted++;
    }
    return o;
}

// Does p[o, hi) look like the start of a BAM record?  Field ranges of the fixed part, the sizes
// adding up inside block_size, the NUL that ends the read name.  `whole` = the record ends
// inside [o, hi) (then *next is where the following one starts).
inline bool plausible_record(const uint8_t* p, size_t o, size_t hi, uint32_t n_refs, size_t* next, bool* whole)
{
    if (hi - o < 36) return false;
    const int32_t bs = rdi32(p + o);
    if (bs < 32 || bs > (1 << 28)) return false;
    const uint8_t* r = p + o + 4;
    const int32_t ref = rdi32(r), pos0 = rdi32(r + 4), l_seq = rdi32(r + 16), nref = rdi32(r + 20), npos = rdi32(r + 24);
    const uint32_t l_name = r[8], n_cig = rd16(r + 12);
    if (ref < -1 || (ref >= 0 && (uint32_t)ref >= n_refs) || nref < -1 || (nref >= 0 && (uint32_t)nref >= n_refs)) return false;
    if (pos0 < -1 || npos < -1 || l_seq < 0 || l_name == 0) return false;
    if (32ull + l_name + 4ull * n_cig + ((uint64_t)l_seq + 1) / 2 + (uint64_t)l_seq > (uint64_t)bs) return false;
    if (o + 36 + l_name <= hi && r[32 + l_name - 1] != 0) return false;
    *next = o + 4 + (size_t)bs;
    *whole = *next <= hi;
    return true;
}

// Where the first record that STARTS in [lo, hi) begins, judged from the bytes alone: the first
// offset from which a chain of plausible records runs (six of them, or up to the end of the
// range).  lo itself in an htslib-written file (records never straddle its BGZF blocks); a few
// bytes further where the writer cuts blocks anywhere (htsjdk).  SIZE_MAX: none found.  Only a
// guess: the caller checks it against where the previous share really stopped.
size_t find_record_start(const uint8_t* p, size_t lo, size_t hi, uint32_t n_refs)
{
    const size_t scan_end = std::min(hi, lo + (1u << 17));
    for (size_t o = lo; o < scan_end; o++) {
        size_t at = o, next = 0;
        bool whole = false, ok = true;
        for (int k = 0; k < 6 && at < hi; k++) {
            if (!plausible_record(p, at, hi, n_refs, &next, &whole)) { ok = hi - at < 36 && k > 0; break; }
            if (!whole) { ok = k > 0; break; }
            at = next;
        }
        if (ok) return o;
    }
    return SIZE_MAX;
}

// Shares of a batch, in file order: their runs become segments of the per-reference lists;
// coordinate order across shares and references is checked here (mscc.pyx:351-366: no reference
// revisited, positions non-decreasing).
struct BamMerger {
    pmio_bam* b;
    std::map<uint32_t, uint64_t> hist;
    std::vector<uint64_t> hist_small = std::vector<uint64_t>(512, 0);
    int32_t cur_ref = -1;
    uint32_t last_pos = 0;
    std::vector<uint8_t> seen;
    std::vector<uint32_t> len_order;    // lengths in file order of first occurrence
    bool sorted = true;

    void take(const pmio_chunk& c, const ShareStats& ss)
    {
        for (uint32_t ql : ss.first_seen) {
            const bool known = ql < hist_small.size() ? hist_small[ql] != 0 : hist.count(ql) != 0;
            if (!known) len_order.push_back(ql);
        }
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
                if (b->run_refs.size() < (1u << 22)) b->run_refs.push_back(r.ref);
                else b->run_overflow = true;
            }
            if (r.first_pos < last_pos) sorted = false;
            last_pos = r.last_pos;
            if (!r.sorted || (b->counts[r.ref] && r.first_pos < b->last_pos[r.ref])) b->ref_ok[r.ref] = 0;
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
    // One streaming pass while the reverse ends come in order (always, when the reads have one
    // length): both lists are written once, nothing else is allocated.
    out.fwd.reserve(n / 2 + n / 8 + 16);
    out.rev.reserve(n / 2 + n / 8 + 16);
    uint32_t last_f = 0, last_e = 0;
    bool ends_sorted = true;
    for (const pmio_bam::Seg& s : segs) {
        const uint32_t* pos = s.c->pos.data();
        const uint32_t* qlen = s.c->qlen.data();
        const uint8_t* rev = s.c->rev.data();
        size_t i = s.begin;
        for (; i < s.end; i++) {
            if (!rev[i]) {
                if (pos[i] == last_f) continue;
                last_f = pos[i];
                out.fwd.push_back(pos[i]);
                out.fwd_len_sum += qlen[i];
            } else {
                const uint32_t e = pos[i] + qlen[i] - 1;
                if (e == last_e) continue;              // that bit is set: the first read's length stays
                if (e < last_e) { ends_sorted = false; break; }
                last_e = e;
                out.rev.push_back(e);
                out.rev_len_sum += qlen[i];
            }
        }
        if (!ends_sorted) break;
    }
    if (ends_sorted) return;
    // Reads of several lengths: the ends are out of order.  Collect them all, sort stably (the
    // first read to set a bit is the one whose length counts), drop repeats.
    out.fwd.clear(); out.rev.clear();
    out.fwd_len_sum = out.rev_len_sum = 0;
    std::vector<uint32_t> ends, lens;
    ends.reserve(n / 2 + n / 8 + 16);
    lens.reserve(n / 2 + n / 8 + 16);
    last_f = 0;
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
                ends.push_back(pos[i] + qlen[i] - 1);
                lens.push_back(qlen[i]);
            }
        }
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
    WorkerPool pool((size_t)threads - 1);
    std::vector<ShareStats> share_stats;
    ShareStats straddler_stats;
    // PMIO_TIMING=1: seconds per stage on stderr (debug aid)
    static const bool timing = getenv("PMIO_TIMING") != nullptr;
    auto now = []() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    double t_decode = 0, t_validate = 0, t_pack = 0, t_mark = now();
    double ts_inflate = 0, ts_filter = 0;       // thread-seconds inside the workers
    // A batch bounds the buffer and is sized to stay in the last-level cache while giving every
    // thread some sixteen shares (measured, 8 threads: 1024-block batches 0.34 s, 128-block ones
    // 0.52 s, 4096-block ones 0.39 s for the same file); inside it the unit of work is a SHARE: a
    // run of whole BGZF blocks of about SHARE_BYTES inflated, small enough that the thread which
    // inflated it filters it while it is still in its own cache.
    size_t BATCH_BLOCKS = std::min<size_t>(4096, std::max<size_t>(512, 128 * (size_t)threads)), SHARE_BYTES = 512u << 10;
    if (const char* e = getenv("PMIO_BATCH_BLOCKS")) BATCH_BLOCKS = std::max<size_t>(1, (size_t)atoll(e));
    if (const char* e = getenv("PMIO_SHARE_BYTES")) SHARE_BYTES = std::max<size_t>(1, (size_t)atoll(e));
    // not a std::vector: nothing may zero (= touch, sequentially) 256 MiB before the workers write it
    struct RawBuffer {
        uint8_t* p = nullptr;
        size_t cap = 0;
        ~RawBuffer() { free(p); }
        void grow(size_t n, size_t keep)
        {
            if (n <= cap) return;
            uint8_t* q = (uint8_t*)malloc(n);
            if (!q) throw std::bad_alloc();
            if (keep) memcpy(q, p, keep);
            free(p);
            p = q;
            cap = n;
        }
    } buf;
    struct Share { size_t blk_lo, blk_hi, lo, hi, start, stop; };
    std::vector<Share> shares;
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
        buf.grow((out_off + (1u << 20)) & ~(size_t)((1u << 20) - 1), carry);
        const char* const inflate_err = "BGZF block does not inflate (corrupt data or CRC mismatch)";
        // ---- the header (first batch): the main thread inflates blocks until it is complete
        size_t used = 0, blk0 = 0;
        while (header.stage != 2 && blk0 < blocks.size()) {
            const Block& k = blocks[blk0++];
            if (!inflate_block(m.p + k.data_off, k.csize, buf.p + k.out_off, k.isize, k.crc)) { g_err = inflate_err; bad = true; break; }
            const long long h = header.feed(buf.p + used, k.out_off + k.isize - used);
            if (h < 0) { bad = true; break; }
            used += (size_t)h;
        }
        if (bad) break;
        if (header.stage != 2) {            // a header longer than the batch: carry what is left of it
            carry = out_off - used;
            if (carry && used) memmove(buf.p, buf.p + used, carry);
            continue;
        }
        // ---- shares: [used, ...) cut at block boundaries every SHARE_BYTES or so
        shares.clear();
        {
            Share s{blk0, blk0, used, used, used, used};
            for (size_t j = blk0; j < blocks.size(); j++) {
                s.blk_hi = j + 1;
                s.hi = blocks[j].out_off + blocks[j].isize;
                if (s.hi - s.lo >= SHARE_BYTES && j + 1 < blocks.size()) {
                    shares.push_back(s);
                    s = Share{j + 1, j + 1, s.hi, s.hi, SIZE_MAX, s.hi};
                }
            }
            s.hi = out_off;
            if (s.hi > s.lo) shares.push_back(s);
        }
        const size_t n_shares = shares.size();
        const uint32_t n_refs = (uint32_t)b->names.size();
        if (share_stats.size() < n_shares) share_stats.resize(n_shares);
        const size_t store0 = b->store.size();      // one chunk per share (the rare record that straddles a share's end gets its own, later)
        for (size_t i = 0; i < n_shares; i++) b->store.emplace_back(new pmio_chunk());
        // ---- every share: inflate, guess where its first record starts, filter -- one thread, one pass.
        //      Share 0 starts at a known record boundary.  The others start at a BGZF block boundary,
        //      which IS a record boundary in an htslib-written file and a few bytes before one
        //      otherwise (find_record_start); a share stops at the record that crosses its end.
        std::atomic<size_t> next_share{0};
        std::atomic<int> failure{0};                // 1 inflate, 2 exception
        std::atomic<uint64_t> ns_inflate{0}, ns_filter{0};
        const uint8_t* p = buf.p;
        auto work = [&]() {
            for (size_t i; (i = next_share.fetch_add(1)) < n_shares && !failure.load(std::memory_order_relaxed);) {
                Share& s = shares[i];
                try {       // an exception must not leave a worker thread
                    const double t0 = timing ? now() : 0;
                    size_t j = s.blk_lo;
                    for (; j + 1 < s.blk_hi; j += 2) {          // in pairs: two decode chains overlap
                        const Block &k0 = blocks[j], &k1 = blocks[j + 1];
                        const uint8_t* const src[2] = {m.p + k0.data_off, m.p + k1.data_off};
                        const size_t csize[2] = {k0.csize, k1.csize}, isize[2] = {k0.isize, k1.isize};
                        uint8_t* const dst[2] = {buf.p + k0.out_off, buf.p + k1.out_off};
                        const uint32_t crc[2] = {k0.crc, k1.crc};
                        if (!inflate_block_pair(src, csize, dst, isize, crc)) { failure.store(1); return; }
                    }
                    for (; j < s.blk_hi; j++) {
                        const Block& k = blocks[j];
                        if (!inflate_block(m.p + k.data_off, k.csize, buf.p + k.out_off, k.isize, k.crc)) { failure.store(1); return; }
                    }
                    const double t1 = timing ? now() : 0;
                    share_stats[i].reset();
                    if (i > 0) s.start = find_record_start(p, s.lo, s.hi, n_refs);
                    s.stop = s.start == SIZE_MAX ? s.lo
                           : parse_share(p, s.start, s.hi, s.hi, n_refs, mapq_criteria, *b->store[store0 + i], share_stats[i]);
                    if (timing) {
                        ns_inflate += (uint64_t)((t1 - t0) * 1e9);
                        ns_filter += (uint64_t)((now() - t1) * 1e9);
                    }
                } catch (...) {
                    failure.store(2);
                }
            }
        };
        pool.run(work, n_shares - 1);
        if (failure.load() == 2) throw std::bad_alloc();
        if (failure.load() == 1) { g_err = inflate_err; bad = true; break; }
        ts_inflate += ns_inflate.load() * 1e-9;
        ts_filter += ns_filter.load() * 1e-9;
        { const double t = now(); t_decode += t - t_mark; t_mark = t; }
        // ---- the guesses are checked in file order: share i+1's start is right iff it is where
        //      share i, and the record that straddled its end, really stopped.  A wrong guess (or
        //      none) is filtered again from the true boundary; a share that lies inside one long
        //      record is dropped.  The record cut by the end of the batch is carried over.
        size_t at = used;
        for (size_t i = 0; i < n_shares && !bad; i++) {
            Share& s = shares[i];
            pmio_chunk& c = *b->store[store0 + i];
            if (at >= s.hi) { c = pmio_chunk(); continue; }
            if (s.start != at) {
                c = pmio_chunk();
                share_stats[i].reset();
                s.stop = parse_share(p, at, s.hi, s.hi, n_refs, mapq_criteria, c, share_stats[i]);
                b->st.reparsed_shares++;
            }
            if (share_stats[i].err) { g_err = share_stats[i].err; bad = true; break; }
            merger.take(c, share_stats[i]);
            at = s.stop;
            if (at >= s.hi) continue;
            // the record starting at `at` crosses the end of the share (or of the batch)
            pmio_chunk one;
            straddler_stats.reset();
            const size_t stop2 = parse_share(p, at, at + 1, out_off, n_refs, mapq_criteria, one, straddler_stats);
            if (straddler_stats.err) { g_err = straddler_stats.err; bad = true; break; }
            if (straddler_stats.incomplete) {       // it runs past the batch: nothing after it can be right
                for (size_t k = i + 1; k < n_shares; k++) *b->store[store0 + k] = pmio_chunk();
                break;
            }
            if (one.pos.empty()) {
                merger.take(one, straddler_stats);          // counters only; no segment points into `one`
            } else {
                b->store.emplace_back(new pmio_chunk(std::move(one)));
                merger.take(*b->store.back(), straddler_stats);
            }
            at = stop2;
        }
        if (bad) break;
        { const double t = now(); t_validate += t - t_mark; t_mark = t; }
        carry = out_off - std::min(at, out_off);
        if (carry && at) memmove(buf.p, buf.p + at, carry);
    }
    if (!bad && header.stage != 2) { g_err = "truncated BAM header"; bad = true; }
    if (!bad && carry) { g_err = "truncated BAM file (incomplete record at the end)"; bad = true; }
    if (bad) return PMIO_ERR_FORMAT;
    b->st.sorted = merger.sorted ? 1 : 0;
    for (size_t l = 0; l < merger.hist_small.size(); l++)
        if (merger.hist_small[l]) merger.hist[(uint32_t)l] += merger.hist_small[l];
    for (uint32_t ql : merger.len_order) { b->hist_len.push_back(ql); b->hist_cnt.push_back(merger.hist[ql]); }
    // packed form per reference, references spread over the threads (longest first)
    {   // (every reference whose own reads are in order, whatever the rest of the file looks like)
        b->packed.resize(b->names.size());
        std::vector<size_t> order;
        for (size_t i = 0; i < b->names.size(); i++)
            if (b->counts[i] && b->ref_ok[i]) order.push_back(i);
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
        if (!order.empty()) pool.run(work, order.size() - 1);
        if (oom.load()) throw std::bad_alloc();
    }
    t_pack = now() - t_mark;
    if (timing)
        fprintf(stderr, "[pmio bam] %d threads, %zu-block batches, %zu-byte shares: inflate + filter %.3f s "
                        "(thread-seconds: inflate + CRC %.2f, filter %.2f), validate + carry %.3f s (%u shares filtered again), pack %.3f s\n", threads, BATCH_BLOCKS,
                SHARE_BYTES, t_decode, ts_inflate, ts_filter, t_validate, b->st.reparsed_shares, t_pack);
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
    if (!b->ref_ok[i]) return fail(PMIO_ERR_ARG, "the reads of this reference are not coordinate sorted");
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

extern "C" int pmio_bam_ref_sorted(const pmio_bam* b, uint32_t i)
{
    if (!b || i >= b->names.size()) return fail(PMIO_ERR_ARG, "reference index out of range");
    return b->ref_ok[i] ? 1 : 0;
}

extern "C" int pmio_bam_ref_order(const pmio_bam* b, const int32_t** refs, uint64_t* n, int* overflow)
{
    if (!b) return fail(PMIO_ERR_ARG, "null argument");
    if (refs) *refs = b->run_refs.data();
    if (n) *n = b->run_refs.size();
    if (overflow) *overflow = b->run_overflow ? 1 : 0;
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
    // pmio_bw_prefetch: every chromosome's intervals at one threshold, by index into `names`
    struct Track {
        std::vector<uint32_t> begins, ends;
        std::vector<float> values;
    };
    std::vector<Track> cache;
    bool have_cache = false;
    float cache_threshold = 0;
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

// One data block (a bedGraph, varStep or fixedStep section of chromosome `id`) appended to B/E/V;
// `buf` is the caller's inflate scratch.  Returns an error message, or nullptr.
static const char* bw_decode_block(const pmio_bw* w, const Leaf& lf, uint32_t id, uint64_t chrom_len, float threshold,
                                   std::vector<uint8_t>& buf, std::vector<uint32_t>& B, std::vector<uint32_t>& E,
                                   std::vector<float>& V)
{
    if (!in_file(w, lf.off, lf.size)) return "BigWig data block outside the file";
    const uint8_t* d = w->m.p + lf.off;
    size_t dn = (size_t)lf.size;
    if (w->uncompress_buf) {
        if (buf.size() < std::max<size_t>(w->uncompress_buf, 64)) buf.resize(std::max<size_t>(w->uncompress_buf, 64));
        // a zlib stream (RFC 1950): 2 header bytes, raw DEFLATE, Adler-32.  The library's own
        // inflate first, vouched for by the checksum; zlib's uncompress for whatever it refuses.
        static const bool zlib_only = getenv("PMIO_ZLIB_ONLY") != nullptr;
        size_t own = 0;
        const bool fast = !zlib_only && dn > 6 && (d[0] & 0x0F) == 8 && ((d[0] << 8 | d[1]) % 31) == 0 && !(d[1] & 0x20) &&
                          own_inflate_upto(d + 2, dn - 6, buf.data(), buf.size(), &own) &&
                          (uint32_t)adler32(adler32(0L, Z_NULL, 0), buf.data(), (uInt)own) ==
                              ((uint32_t)d[dn - 4] << 24 | (uint32_t)d[dn - 3] << 16 | (uint32_t)d[dn - 2] << 8 | d[dn - 1]);
        uLongf got = fast ? (uLongf)own : (uLongf)buf.size();
        int zr = fast ? Z_OK : uncompress(buf.data(), &got, d, (uLong)lf.size);
        while (zr == Z_BUF_ERROR && buf.size() < (1u << 28)) {       // header understated the size
            buf.resize(buf.size() * 2);
            got = (uLongf)buf.size();
            zr = uncompress(buf.data(), &got, d, (uLong)lf.size);
        }
        if (zr != Z_OK) return "BigWig data block does not inflate";
        d = buf.data();
        dn = (size_t)got;
    }
    if (dn < 24) return "short BigWig data block";
    if (rd32(d) != id) return nullptr;
    const uint32_t start0 = rd32(d + 4), step = rd32(d + 12), span = rd32(d + 16);
    const uint32_t type = d[20], count = rd16(d + 22);
    const size_t item = type == 1 ? 12 : type == 2 ? 8 : 4;
    if (type < 1 || type > 3 || dn < 24 + count * item) return "corrupt BigWig data block";
    for (uint32_t i = 0; i < count; i++) {
        const uint8_t* it = d + 24 + i * item;
        uint32_t s, e;
        float v;
        if (type == 1) { s = rd32(it); e = rd32(it + 4); v = rdf32(it + 8); }
        else if (type == 2) { s = rd32(it); e = s + span; v = rdf32(it + 4); }
        else { s = start0 + i * step; e = s + span; v = rdf32(it); }
        if (s >= chrom_len || e == 0) continue;                  // intervals(chrom, 0, size): overlap only
        if (threshold > 0 && !(v >= threshold)) continue;        // reader/bigwig.pyx:174-175
        B.push_back(s); E.push_back(e); V.push_back(v);
    }
    return nullptr;
}

static int bw_intervals_impl(pmio_bw* w, const char* chrom, float threshold, const uint32_t** begins,
                             const uint32_t** ends, const float** values, uint64_t* n)
{
    size_t ci = 0;
    for (; ci < w->names.size(); ci++)
        if (w->names[ci] == chrom) break;
    if (ci == w->names.size()) return fail(PMIO_ERR_KEY, chrom);
    if (w->have_cache && w->cache_threshold == threshold) {      // decoded by pmio_bw_prefetch
        const pmio_bw::Track& t = w->cache[ci];
        if (begins) *begins = t.begins.data();
        if (ends) *ends = t.ends.data();
        if (values) *values = t.values.data();
        if (n) *n = t.begins.size();
        return PMIO_OK;
    }
    std::vector<Leaf> leaves;
    if (!walk_rtree(w, w->index_off + 48, w->ids[ci], leaves, 0)) return fail(PMIO_ERR_FORMAT, "corrupt BigWig index");
    w->begins.clear(); w->ends.clear(); w->values.clear();
    std::vector<uint8_t> buf;
    for (const Leaf& lf : leaves)
        if (const char* err = bw_decode_block(w, lf, w->ids[ci], w->lens[ci], threshold, buf, w->begins, w->ends, w->values))
            return fail(PMIO_ERR_FORMAT, err);
    if (begins) *begins = w->begins.data();
    if (ends) *ends = w->ends.data();
    if (values) *values = w->values.data();
    if (n) *n = w->begins.size();
    return PMIO_OK;
}

// Every chromosome at once: the data blocks of the whole file are decoded by a pool of threads
// (runs of 32 consecutive blocks of a chromosome per task), then stitched per chromosome.
static int bw_prefetch_impl(pmio_bw* w, float threshold, int threads)
{
    if (threads <= 0) threads = (int)std::max(1u, std::thread::hardware_concurrency());
    threads = std::min(threads, 64);
    struct Piece { size_t ci; size_t lo, hi; pmio_bw::Track out; };
    auto now = []() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    const double t0 = now();
    const size_t nc = w->names.size();
    std::vector<std::vector<Leaf>> leaves(nc);
    std::vector<Piece> pieces;
    for (size_t ci = 0; ci < nc; ci++) {
        if (!walk_rtree(w, w->index_off + 48, w->ids[ci], leaves[ci], 0)) return fail(PMIO_ERR_FORMAT, "corrupt BigWig index");
        for (size_t lo = 0; lo < leaves[ci].size(); lo += 32)
            pieces.push_back(Piece{ci, lo, std::min(lo + 32, leaves[ci].size()), {}});
    }
    std::atomic<size_t> next{0};
    std::atomic<const char*> error{nullptr};
    std::atomic<bool> oom{false};
    auto decode = [&]() {
        std::vector<uint8_t> buf;
        try {
            for (size_t k; (k = next.fetch_add(1)) < pieces.size() && !error.load();) {
                Piece& pc = pieces[k];
                pmio_bw::Track out;     // on this thread's stack: neighbouring pieces belong to other threads
                for (size_t j = pc.lo; j < pc.hi; j++)
                    if (const char* err = bw_decode_block(w, leaves[pc.ci][j], w->ids[pc.ci], w->lens[pc.ci], threshold, buf,
                                                          out.begins, out.ends, out.values)) {
                        error.store(err);
                        return;
                    }
                pc.out = std::move(out);
            }
        } catch (...) {
            oom.store(true);
        }
    };
    const double t1 = now();
    WorkerPool pool((size_t)threads - 1);
    if (!pieces.empty()) pool.run(decode, pieces.size() - 1);
    const double t2 = now();
    if (oom.load()) throw std::bad_alloc();
    if (error.load()) return fail(PMIO_ERR_FORMAT, error.load());
    // pieces are in (chromosome, file) order
    std::vector<size_t> first(nc, pieces.size());       // a chromosome without blocks: no pieces
    for (size_t k = pieces.size(); k-- > 0;) first[pieces[k].ci] = k;
    std::vector<pmio_bw::Track> cache(nc);
    std::atomic<size_t> next_chrom{0};
    auto stitch = [&]() {
        try {
            for (size_t ci; (ci = next_chrom.fetch_add(1)) < nc;) {
                size_t total = 0;
                for (size_t k = first[ci]; k < pieces.size() && pieces[k].ci == ci; k++) total += pieces[k].out.begins.size();
                pmio_bw::Track& t = cache[ci];
                t.begins.reserve(total); t.ends.reserve(total); t.values.reserve(total);
                for (size_t k = first[ci]; k < pieces.size() && pieces[k].ci == ci; k++) {
                    const pmio_bw::Track& o = pieces[k].out;
                    t.begins.insert(t.begins.end(), o.begins.begin(), o.begins.end());
                    t.ends.insert(t.ends.end(), o.ends.begin(), o.ends.end());
                    t.values.insert(t.values.end(), o.values.begin(), o.values.end());
                }
            }
        } catch (...) {
            oom.store(true);
        }
    };
    if (nc) pool.run(stitch, nc - 1);
    if (oom.load()) throw std::bad_alloc();
    if (getenv("PMIO_TIMING"))
        fprintf(stderr, "[pmio bigwig] %d threads, %zu pieces: index %.4f s, decode %.4f s, stitch %.4f s\n", threads,
                pieces.size(), t1 - t0, t2 - t1, now() - t2);
    w->cache.swap(cache);
    w->cache_threshold = threshold;
    w->have_cache = true;
    return PMIO_OK;
}

extern "C" int pmio_bw_prefetch(pmio_bw* w, float threshold, int threads)
{
    if (!w) return fail(PMIO_ERR_ARG, "null argument");
    return guarded([&]() { return bw_prefetch_impl(w, threshold, threads); });
}
