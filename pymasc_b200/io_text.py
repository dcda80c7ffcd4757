"""Host-side text readers that hand packed arrays to the calculator.

BAM/SAM decoding and BigWig parsing stay on the host (north_star); pysam and pyBigWig
are not installable in this image, so the SAM and bedGraph text twins of the formats
are read here.  The filter and coordinate rules are the reference's:

* skip read2, MAPQ < threshold, unmapped, duplicate   PyMaSC/handler/read.py:62-90
* pos = reference_start + 1 (SAM POS is already 1-based) PyMaSC/handler/read.py:138
* readlen = query length from the CIGAR (M/I/S/=/X)      PyMaSC/handler/read.py:139
* mappable intervals: value >= 1.0, 0-based half-open     PyMaSC/reader/bigwig.pyx:171-177
"""
import re

import numpy as np

_CIGAR = re.compile(r"(\d+)([MIDNSHP=X])")
_QUERY_OPS = frozenset("MIS=X")


def query_length(cigar):
    """pysam's infer_query_length(): sum of M/I/S/=/X operations."""
    return sum(int(n) for n, op in _CIGAR.findall(cigar) if op in _QUERY_OPS)


def read_sam(path, mapq_criteria=1, with_readlen_counter=False):
    """Parse a SAM file.

    Returns (references, lengths, reads) where reads maps chrom -> dict of numpy arrays
    ``pos`` (1-based leftmost base, int64), ``readlen`` (int64), ``is_reverse`` (bool),
    in file order (coordinate sorted input stays sorted).  With ``with_readlen_counter`` a
    fourth value: the query-length counter of estimate_readlen (core/readlen.pyx:150-156:
    mapped, not duplicate, MAPQ >= criteria, with a reference; read2 included).
    """
    refs, lens = [], []
    acc = {}
    counter = {}
    with open(path) as fh:
        for line in fh:
            if line.startswith("@"):
                if line.startswith("@SQ"):
                    tags = dict(f.split(":", 1) for f in line.rstrip("\n").split("\t")[1:])
                    refs.append(tags["SN"])
                    lens.append(int(tags["LN"]))
                continue
            f = line.split("\t")
            flag = int(f[1])
            if f[2] == "*":
                continue
            good = not (flag & 0x4 or flag & 0x400 or int(f[4]) < mapq_criteria)
            qlen = query_length(f[5]) if good else 0
            if good:
                counter[qlen] = counter.get(qlen, 0) + 1
            if flag & 0x80 or not good or qlen == 0:
                continue
            rec = acc.setdefault(f[2], ([], [], []))
            rec[0].append(int(f[3]))
            rec[1].append(qlen)
            rec[2].append(bool(flag & 0x10))
    reads = {c: {"pos": np.asarray(p, np.int64), "readlen": np.asarray(l, np.int64),
                 "is_reverse": np.asarray(r, bool)} for c, (p, l, r) in acc.items()}
    if with_readlen_counter:
        return refs, lens, reads, counter
    return refs, lens, reads


def read_bedgraph(path, threshold=1.0):
    """chrom -> (begins, ends) uint32 arrays of intervals with value >= threshold."""
    acc = {}
    with open(path) as fh:
        for line in fh:
            if not line.strip() or line.startswith(("#", "track", "browser")):
                continue
            c, b, e, v = line.split()[:4]
            if float(v) >= threshold:
                rec = acc.setdefault(c, ([], []))
                rec[0].append(int(b))
                rec[1].append(int(e))
    return {c: (np.asarray(b, np.uint32), np.asarray(e, np.uint32)) for c, (b, e) in acc.items()}


def read_chrom_sizes(path):
    out = {}
    with open(path) as fh:
        for line in fh:
            p = line.split()
            if len(p) >= 2:
                out[p[0]] = int(p[1])
    return out
