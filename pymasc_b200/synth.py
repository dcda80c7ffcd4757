"""Synthetic hg38-shaped workloads as packed arrays.

There is no network for real data, so the benchmark and the full-size parity tests use
seeded synthetic reads and mappability with the real hg38 chromosome lengths.  Everything
is generated on the host with numpy; nothing here is on the timed path.
"""
import numpy as np

HG38 = [
    ("chr1", 248956422), ("chr2", 242193529), ("chr3", 198295559), ("chr4", 190214555),
    ("chr5", 181538259), ("chr6", 170805979), ("chr7", 159345973), ("chr8", 145138636),
    ("chr9", 138394717), ("chr10", 133797422), ("chr11", 135086622), ("chr12", 133275309),
    ("chr13", 114364328), ("chr14", 107043718), ("chr15", 101991189), ("chr16", 90338345),
    ("chr17", 83257441), ("chr18", 80373285), ("chr19", 58617616), ("chr20", 64444167),
    ("chr21", 46709983), ("chr22", 50818468), ("chrX", 156040895), ("chrY", 57227415),
]


def genome(scale=1.0, chroms=None):
    """(names, lengths); scale < 1 shrinks every chromosome (tests)."""
    sel = [g for g in HG38 if chroms is None or g[0] in chroms]
    return [n for n, _ in sel], [max(2000, int(l * scale)) for _, l in sel]


def chip_reads(chrom_index, chrom_len, n_reads, read_len=36, seed=20260101,
               peak_fraction=0.3, frag_mean=200.0, frag_sd=30.0):
    """ChIP-like single-end reads of one chromosome.

    70 % background (uniform starts), 30 % in peaks (fragment length ~ N(200, 30^2); the
    forward read starts at c - frag/2, the reverse read's 5' end sits at c + frag/2), strand
    ~ Bernoulli(0.5).  Returns sorted unique (fwd_pos, rev_pos) uint32 arrays: 1-based forward
    5' ends and reverse 5' ends (= leftmost + read_len - 1), i.e. exactly the bits the
    calculator sets.
    """
    rng = np.random.default_rng(seed + chrom_index)
    n_peak = int(n_reads * peak_fraction)
    n_bg = n_reads - n_peak
    n_centres = max(1, int(200000 * chrom_len / 3.1e9))
    centres = rng.integers(1000, max(1001, chrom_len - 1000), size=n_centres)
    c = centres[rng.integers(0, n_centres, size=n_peak)]
    frag = np.clip(rng.normal(frag_mean, frag_sd, size=n_peak), read_len, None).astype(np.int64)
    strand_peak = rng.random(n_peak) < 0.5
    fwd_peak = (c - frag // 2)[~strand_peak]
    rev_peak = (c + frag // 2)[strand_peak]
    strand_bg = rng.random(n_bg) < 0.5
    start_bg = rng.integers(1, max(2, chrom_len - read_len), size=n_bg)
    fwd = np.concatenate([fwd_peak, start_bg[~strand_bg]])
    rev = np.concatenate([rev_peak, start_bg[strand_bg] + read_len - 1])
    fwd = np.unique(np.clip(fwd, 1, chrom_len - 1)).astype(np.uint32)
    rev = np.unique(np.clip(rev, read_len, chrom_len + read_len - 2)).astype(np.uint32)
    return fwd, rev


def mappability(chrom_index, chrom_len, seed=777, mean_run=1500.0, mean_gap=400.0):
    """Alternating mappable / unmappable runs (~79 % mappable): (begins, ends) uint32,
    0-based half-open, sorted and disjoint -- what BigWigReader.fetch(1.0, chrom) yields."""
    rng = np.random.default_rng(seed + chrom_index)
    n = int(chrom_len / (mean_run + mean_gap) * 1.2) + 16
    runs = 1 + rng.geometric(1.0 / mean_run, size=n)
    gaps = 1 + rng.geometric(1.0 / mean_gap, size=n)
    ends = np.cumsum(runs + gaps)
    begins = ends - runs
    keep = ends <= chrom_len
    return begins[keep].astype(np.uint32), ends[keep].astype(np.uint32)


def workload(n_reads=20_000_000, read_len=36, scale=1.0, chroms=None, with_mappability=True,
             seed=20260101):
    """Packed inputs of a whole genome: list of dicts (name, length, fwd, rev, begins, ends)."""
    names, lengths = genome(scale, chroms)
    total = float(sum(lengths))
    out = []
    for name, L in zip(names, lengths):
        idx = [n for n, _ in HG38].index(name)
        n = max(8, int(round(n_reads * L / total)))
        fwd, rev = chip_reads(idx, L, n, read_len, seed)
        b = e = None
        if with_mappability:
            b, e = mappability(idx, L)
        out.append({"name": name, "length": L, "fwd": fwd, "rev": rev, "begins": b, "ends": e})
    return out


def evals(lengths, read_len, max_shift):
    """bit-position x shift evaluations of a run: sum_chrom Lbits x (max_shift+1)."""
    return sum(int(L) + read_len + max_shift + 100 for L in lengths) * (max_shift + 1)
