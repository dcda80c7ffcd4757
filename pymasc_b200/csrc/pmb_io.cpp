// pmb_io.cpp -- libpymasc_io.so (include/pymasc_io.h): BAM (BGZF) and BigWig decoding on the
// host into packed arrays.  zlib + std::thread; no CUDA here.
//
// BAM: the file is memory-mapped, BGZF block headers are walked sequentially (18 bytes per
// <= 64 KiB block), batches of blocks are inflated in parallel straight into their final
// offsets of one buffer (ISIZE in each block's footer gives the offsets up front), and the
// records of the batch are filtered and packed by one pass over that buffer; a record cut by
// the batch boundary is carried into the next batch.
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
#include <map>
#include <string>
#include <thread>
#include <vector>

#include "../../include/pymasc_io.h"

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

// raw-deflate payload -> dst (exactly `isize` bytes); crc checked
bool inflate_block(const uint8_t* src, size_t csize, uint8_t* dst, size_t isize, uint32_t crc)
{
    if (isize == 0) return true;
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
    return ok && (uint32_t)crc32(crc32(0L, Z_NULL, 0), dst, (uInt)isize) == crc;
}

struct Block { size_t data_off; uint32_t csize, isize, crc; size_t out_off; };

}  // namespace

// ====================================================================== BAM
struct pmio_bam {
    std::vector<std::string> names;
    std::vector<uint64_t> lens;
    std::vector<std::vector<uint32_t>> pos, qlen;
    std::vector<std::vector<uint8_t>> rev;
    std::vector<uint32_t> hist_len;
    std::vector<uint64_t> hist_cnt;
    pmio_bam_stats st{};
};

namespace {

struct BamParser {
    pmio_bam* b;
    int mapq;
    // header state
    int stage = 0;              // 0 magic + text, 1 references, 2 records
    uint32_t refs_left = 0;
    std::map<uint32_t, uint64_t> hist;
    std::vector<uint64_t> hist_small = std::vector<uint64_t>(1024, 0);
    int32_t cur_ref = -1;
    uint32_t last_pos = 0;
    std::vector<uint8_t> seen;
    bool sorted = true;

    // consumes as many complete items as [p, p+n) holds; returns bytes consumed, or -1
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
            b->names.reserve(refs_left);
            stage = 1;
        }
        while (stage == 1) {
            if (refs_left == 0) {
                const size_t nr = b->names.size();
                b->pos.resize(nr); b->qlen.resize(nr); b->rev.resize(nr);
                seen.assign(nr, 0);
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
        pmio_bam_stats& st = b->st;
        const uint32_t n_refs = (uint32_t)b->names.size();
        while (n - o >= 4) {
            const int32_t bs = rdi32(p + o);
            if (bs < 32) { g_err = "corrupt BAM record (block_size < 32)"; return -1; }
            if (n - o < 4 + (size_t)bs) break;
            const uint8_t* r = p + o + 4;
            o += 4 + (size_t)bs;
            st.n_records++;
            const int32_t ref = rdi32(r);
            if (ref < 0) continue;                                   // reference_name is None
            if ((uint32_t)ref >= n_refs) { g_err = "BAM record names an unknown reference"; return -1; }
            const int32_t pos0 = rdi32(r + 4);
            const uint32_t l_name = r[8], mq = r[9];
            const uint32_t n_cig = rd16(r + 12), flag = rd16(r + 14);
            st.n_with_ref++;
            if (flag & 0x1) { st.n_paired++; if (flag & 0x80) st.n_read2++; }
            if (flag & 0x4) st.n_unmapped++;
            if (32 + l_name + 4ull * n_cig > (uint32_t)bs) { g_err = "corrupt BAM record (CIGAR overruns)"; return -1; }
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
                if (ql < hist_small.size()) hist_small[ql]++;
                else hist[ql]++;
            }
            if ((flag & 0x80) || !good || ql == 0) continue;        // handler/read.py:62-90, :132-143
            if (ref != cur_ref) {
                if (seen[ref]) sorted = false;
                seen[ref] = 1;
                cur_ref = ref;
                last_pos = 0;
            }
            const uint32_t p1 = (uint32_t)(pos0 + 1);
            if (p1 < last_pos) sorted = false;
            last_pos = p1;
            b->pos[ref].push_back(p1);
            b->qlen[ref].push_back(ql);
            b->rev[ref].push_back((flag & 0x10) ? 1 : 0);
            st.n_accepted++;
        }
        return (long long)o;
    }
};

}  // namespace

extern "C" int pmio_bam_load(const char* path, int mapq_criteria, int threads, pmio_bam** out)
{
    if (!path || !out) return fail(PMIO_ERR_ARG, "null argument");
    *out = nullptr;
    Mapping m;
    int rc = m.open(path);
    if (rc != PMIO_OK) return rc;
    if (threads <= 0) threads = (int)std::max(1u, std::thread::hardware_concurrency());
    threads = std::min(threads, 64);

    pmio_bam* b = new (std::nothrow) pmio_bam();
    if (!b) return fail(PMIO_ERR_NOMEM, "out of memory");
    BamParser parser{b, mapq_criteria};
    const size_t BATCH_BLOCKS = 2048;                 // <= 128 MiB inflated per batch
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
        // ---- parse; keep the cut record for the next batch
        const long long used = parser.feed(buf.data(), out_off);
        if (used < 0) { bad = true; break; }
        carry = out_off - (size_t)used;
        if (carry && used) memmove(buf.data(), buf.data() + used, carry);
    }
    if (!bad && parser.stage != 2) { g_err = "truncated BAM header"; bad = true; }
    if (!bad && carry) { g_err = "truncated BAM file (incomplete record at the end)"; bad = true; }
    if (bad) { delete b; return PMIO_ERR_FORMAT; }
    b->st.sorted = parser.sorted ? 1 : 0;
    for (size_t l = 0; l < parser.hist_small.size(); l++)
        if (parser.hist_small[l]) parser.hist[(uint32_t)l] += parser.hist_small[l];
    for (auto& kv : parser.hist) { b->hist_len.push_back(kv.first); b->hist_cnt.push_back(kv.second); }
    *out = b;
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

extern "C" int pmio_bam_reads(const pmio_bam* b, uint32_t i, const uint32_t** pos, const uint32_t** qlen,
                              const uint8_t** is_reverse, uint64_t* n)
{
    if (!b || i >= b->names.size()) return fail(PMIO_ERR_ARG, "reference index out of range");
    if (pos) *pos = b->pos[i].data();
    if (qlen) *qlen = b->qlen[i].data();
    if (is_reverse) *is_reverse = b->rev[i].data();
    if (n) *n = b->pos[i].size();
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

extern "C" int pmio_bw_open(const char* path, pmio_bw** out)
{
    if (!path || !out) return fail(PMIO_ERR_ARG, "null argument");
    *out = nullptr;
    pmio_bw* w = new (std::nothrow) pmio_bw();
    if (!w) return fail(PMIO_ERR_NOMEM, "out of memory");
    int rc = w->m.open(path);
    if (rc != PMIO_OK) { delete w; return rc; }
    const uint8_t* p = w->m.p;
    if (w->m.n < 64 || rd32(p) != 0x888FFC26u) {
        delete w;
        return fail(PMIO_ERR_FORMAT, std::string(path) + " is not a little-endian BigWig file (bad magic)");
    }
    const uint64_t chrom_tree = rd64(p + 8);
    w->index_off = rd64(p + 24);
    w->uncompress_buf = rd32(p + 52);
    if (!in_file(w, chrom_tree, 32) || rd32(p + chrom_tree) != 0x78CA8C91u ||
        !in_file(w, w->index_off, 48) || rd32(p + w->index_off) != 0x2468ACE0u) {
        delete w;
        return fail(PMIO_ERR_FORMAT, "corrupt BigWig file (chromosome tree or index)");
    }
    const uint32_t key_size = rd32(p + chrom_tree + 8), val_size = rd32(p + chrom_tree + 12);
    if (val_size != 8 || key_size == 0 || key_size > 4096 || !walk_chrom_tree(w, chrom_tree + 32, key_size, 0)) {
        delete w;
        return fail(PMIO_ERR_FORMAT, "corrupt BigWig chromosome tree");
    }
    *out = w;
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

extern "C" int pmio_bw_intervals(pmio_bw* w, const char* chrom, float threshold, const uint32_t** begins,
                                 const uint32_t** ends, const float** values, uint64_t* n)
{
    if (!w || !chrom) return fail(PMIO_ERR_ARG, "null argument");
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
