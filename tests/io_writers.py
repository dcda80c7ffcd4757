"""Minimal BAM (BGZF) and BigWig WRITERS for the reader tests: they produce spec-conformant
files with shapes the reference fixtures lack (records cut by BGZF block boundaries, paired /
unmapped / duplicate flags, reads without CIGAR, several chromosomes, varStep / fixedStep
sections, a two-level R-tree, uncompressed data).  Test infrastructure only."""
import struct
import zlib

CIGAR_OPS = "MIDNSHP=X"


def bgzf_block(payload):
    assert len(payload) <= 65280
    co = zlib.compressobj(6, zlib.DEFLATED, -15)
    data = co.compress(payload) + co.flush()
    bsize = 12 + 6 + len(data) + 8 - 1
    head = struct.pack("<BBBBIBBH", 31, 139, 8, 4, 0, 0, 255, 6) + b"BC" + struct.pack("<HH", 2, bsize)
    return head + data + struct.pack("<II", zlib.crc32(payload) & 0xFFFFFFFF, len(payload))


BGZF_EOF = bgzf_block(b"")


def bam_record(ref_id, pos0, mapq, flag, cigar, name=b"r", l_seq=None):
    """cigar: list of (length, op char).  pos0 is 0-based."""
    nm = name + b"\0"
    cg = b"".join(struct.pack("<I", (n << 4) | CIGAR_OPS.index(op)) for n, op in cigar)
    if l_seq is None:
        l_seq = sum(n for n, op in cigar if op in "MIS=X")
    seq = b"\x11" * ((l_seq + 1) // 2)
    qual = b"\xff" * l_seq
    core = struct.pack("<iiBBHHHiiii", ref_id, pos0, len(nm), mapq, 4680, len(cigar), flag, l_seq, -1, -1, 0)
    body = core + nm + cg + seq + qual
    return struct.pack("<i", len(body)) + body


def write_bam(path, refs, records, block_payload=65280, text=b"@HD\tVN:1.6\tSO:coordinate\n",
              align_records=False):
    """refs: [(name, length)]; records: bytes objects from bam_record.  The uncompressed stream
    is cut into blocks of `block_payload` bytes regardless of record boundaries; with
    `align_records` the way htslib does it: the header in blocks of its own, and a record that
    would not fit starts a new block (no record straddles two blocks unless it is longer than
    one)."""
    head = b"BAM\1" + struct.pack("<i", len(text)) + text + struct.pack("<i", len(refs))
    for name, length in refs:
        nm = name.encode() + b"\0"
        head += struct.pack("<i", len(nm)) + nm + struct.pack("<i", length)
    with open(path, "wb") as fh:
        if not align_records:
            stream = head + b"".join(records)
            for o in range(0, len(stream), block_payload):
                fh.write(bgzf_block(stream[o:o + block_payload]))
        else:
            for o in range(0, len(head), block_payload):
                fh.write(bgzf_block(head[o:o + block_payload]))
            cur = b""
            for r in records:
                if cur and len(cur) + len(r) > block_payload:
                    fh.write(bgzf_block(cur))
                    cur = b""
                cur += r
                while len(cur) > block_payload:           # a record longer than a block
                    fh.write(bgzf_block(cur[:block_payload]))
                    cur = cur[block_payload:]
            if cur:
                fh.write(bgzf_block(cur))
        fh.write(BGZF_EOF)


# ------------------------------------------------------------------------------------- BigWig
def write_bigwig(path, chroms, sections, compress=True, two_level=False):
    """chroms: [(name, length)] (ids in this order); sections: list of dicts
    {chrom_id, type: 'bedgraph'|'varstep'|'fixedstep', items, start, step, span}:
      bedgraph  items = [(start, end, value)]
      varstep   items = [(start, value)], span
      fixedstep items = [value], start, step, span
    Sections must be sorted by (chrom_id, start)."""
    blocks = []
    for s in sections:
        t = s["type"]
        if t == "bedgraph":
            lo, hi = s["items"][0][0], s["items"][-1][1]
            body = b"".join(struct.pack("<IIf", a, b, v) for a, b, v in s["items"])
            head = struct.pack("<IIIIIBBH", s["chrom_id"], lo, hi, 0, 0, 1, 0, len(s["items"]))
        elif t == "varstep":
            lo, hi = s["items"][0][0], s["items"][-1][0] + s["span"]
            body = b"".join(struct.pack("<If", a, v) for a, v in s["items"])
            head = struct.pack("<IIIIIBBH", s["chrom_id"], lo, hi, 0, s["span"], 2, 0, len(s["items"]))
        else:
            lo = s["start"]
            hi = lo + (len(s["items"]) - 1) * s["step"] + s["span"]
            body = b"".join(struct.pack("<f", v) for v in s["items"])
            head = struct.pack("<IIIIIBBH", s["chrom_id"], lo, hi, s["step"], s["span"], 3, 0, len(s["items"]))
        raw = head + body
        blocks.append((s["chrom_id"], lo, hi, zlib.compress(raw) if compress else raw, len(raw)))

    key_size = max(len(n) for n, _ in chroms) + 1
    header_size = 64
    chrom_tree_off = header_size
    tree = struct.pack("<IIIIQQ", 0x78CA8C91, len(chroms), key_size, 8, len(chroms), 0)
    tree += struct.pack("<BBH", 1, 0, len(chroms))
    for i, (n, ln) in sorted(enumerate(chroms), key=lambda x: x[1][0]):
        tree += n.encode().ljust(key_size, b"\0") + struct.pack("<II", i, ln)
    data_off = chrom_tree_off + len(tree)
    data = struct.pack("<Q", len(blocks))
    offs = []
    for _, _, _, payload, _ in blocks:
        offs.append(data_off + len(data))
        data += payload
    index_off = data_off + len(data)

    def leaf_node(idx):
        out = struct.pack("<BBH", 1, 0, len(idx))
        for i in idx:
            c, lo, hi, payload, _ = blocks[i]
            out += struct.pack("<IIIIQQ", c, lo, c, hi, offs[i], len(payload))
        return out

    rhead_size = 48
    if two_level and len(blocks) > 1:
        half = len(blocks) // 2
        groups = [list(range(0, half)), list(range(half, len(blocks)))]
        root_size = 4 + 24 * len(groups)
        leaves, child_offs, o = [], [], index_off + rhead_size + root_size
        for g in groups:
            node = leaf_node(g)
            leaves.append(node)
            child_offs.append(o)
            o += len(node)
        root = struct.pack("<BBH", 0, 0, len(groups))
        for g, co in zip(groups, child_offs):
            root += struct.pack("<IIIIQ", blocks[g[0]][0], blocks[g[0]][1], blocks[g[-1]][0], blocks[g[-1]][2], co)
        nodes = root + b"".join(leaves)
    else:
        nodes = leaf_node(list(range(len(blocks))))
    rhead = struct.pack("<IIQIIIIQII", 0x2468ACE0, 256, len(blocks), blocks[0][0], blocks[0][1],
                        blocks[-1][0], blocks[-1][2], index_off, 1, 0)
    max_raw = max(b[4] for b in blocks)
    header = struct.pack("<IHHQQQHHQQIQ", 0x888FFC26, 4, 0, chrom_tree_off, data_off, index_off, 0, 0, 0, 0,
                         max_raw if compress else 0, 0)
    assert len(header) == 64 and len(rhead) == 48
    with open(path, "wb") as fh:
        fh.write(header + tree + data + rhead + nodes)
