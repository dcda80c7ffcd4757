"""libpymasc_io.so (include/pymasc_io.h) and pymasc_b200.io_native: native BAM / BigWig decoding.

Pins: the reference's own binary fixtures (tests/golden/ENCFF000RMB-test.bam and
hg19_36mer-test.bigwig, copied byte for byte from the reference's tests/data by
tests/golden/make_golden.py) must decode to exactly the arrays parsed from their text twins
(.sam / .bedGraph -> fixture_encff000rmb.npz); synthetic files written by tests/io_writers.py
cover what the fixtures lack.  No GPU needed.
"""
import ctypes
import os
import struct

import numpy as np
import pytest

import helpers
import io_writers as W
from oracle_engine import OracleEngine
from pymasc_b200 import CCBitArrayCalculator, ReadUnsortedError, io_native

BAM = os.path.join(helpers.GOLDEN, "ENCFF000RMB-test.bam")
BIGWIG = os.path.join(helpers.GOLDEN, "hg19_36mer-test.bigwig")


def test_library_exports_every_declared_symbol():
    L = io_native.load_library()
    header = open(os.path.join(helpers.ROOT, "include", "pymasc_io.h")).read()
    import re
    declared = set(re.findall(r"\b(pmio_[a-z0-9_]+)\s*\(", header))
    assert declared == set(io_native.EXPORTS)
    for name in declared:
        assert hasattr(L, name), name
    assert ctypes.sizeof(io_native.BamStats) == 56


# ---------------------------------------------------------------------------------------- BAM
@pytest.mark.parametrize("threads", [1, 4])
def test_reference_bam_fixture_decodes_to_the_sam_arrays(threads):
    fx = helpers.load_fixture()
    bam = io_native.BamFile(BAM, mapq_criteria=10, threads=threads)
    assert list(bam.references) == fx["references"] and list(bam.lengths) == fx["lengths"]
    assert list(bam.reads) == ["chr1"] and bam.sorted
    r = bam.reads["chr1"]
    assert np.array_equal(r["pos"], fx["pos"])
    assert np.array_equal(r["readlen"], fx["readlen"])
    assert np.array_equal(r["is_reverse"], fx["is_reverse"])
    assert bam.stats["n_records"] == 2501 and bam.stats["n_accepted"] == 1292
    # read-length estimate (core/readlen.pyx): the fixture is a 36-mer library
    assert io_native.estimate_readlen(bam, "MEDIAN", 10) == 36
    assert io_native.estimate_readlen(bam, "MODE", 10) == 36
    assert io_native.estimate_readlen(bam, "MIN", 10) == min(bam.readlen_counter)
    assert io_native.estimate_readlen(bam, "MAX", 10) == max(bam.readlen_counter)
    c = bam.readlen_counter
    assert io_native.estimate_readlen(bam, "MEAN", 10) == int(round(sum(k * v for k, v in c.items()) / sum(c.values())))


def test_bam_to_golden_tables_through_the_calculator():
    """BAM + BigWig files in, the reference's golden NCC/MSCC curves out (configs[0] from the
    binary files; the device is replaced by the oracle engine in the CPU suite)."""
    gold = helpers.load_golden_tables()
    bam = io_native.BamFile(BAM, mapq_criteria=10)
    bw = io_native.BigWigReader(BIGWIG)
    calc = CCBitArrayCalculator(300, 36, bam.references, bam.lengths, bw, engine=OracleEngine())
    io_native.feed_bam(calc, bam)
    calc.finishup_calculation()
    whole = calc.get_whole_result()
    bw.close()
    for got, want in ((whole.chroms["chr1"].cc, gold["cc"]["chr1"]),
                      (whole.mappable_chroms["chr1"].cc, gold["mscc"]["chr1"])):
        want = np.array(want)
        assert np.max(np.abs(np.asarray(got) - want) / np.abs(want)) <= 1e-9


def synthetic_records():
    """(records, expected accepted reads per chromosome, expected read-length counter)."""
    rng = np.random.default_rng(5)
    recs, want, hist = [], {"c1": [], "c2": [], "c3": []}, {}
    pos = 10
    for ref, chrom, n in ((0, "c1", 3000), (2, "c3", 1500)):
        pos = 10
        for i in range(n):
            pos += int(rng.integers(0, 40))
            kind = i % 11
            flag, mapq = (16 if rng.random() < 0.5 else 0), int(rng.integers(0, 60))
            cigar = [(int(rng.integers(20, 60)), "M")]
            if kind == 1:
                cigar = [(5, "S"), (30, "M"), (2, "I"), (3, "D"), (10, "="), (100, "N"), (4, "X"), (6, "H")]
            if kind == 2:
                flag |= 0x4                       # unmapped
            if kind == 3:
                flag |= 0x400                     # duplicate
            if kind == 4:
                flag |= 0x1 | 0x80                # paired, read2
            if kind == 5:
                flag |= 0x1 | 0x40                # paired, read1
            if kind == 6:
                cigar = []                        # no CIGAR: infer_query_length() is None
            name = b"read%07d" % i + b"x" * int(rng.integers(0, 30))
            recs.append(W.bam_record(ref, pos - 1, mapq, flag, cigar, name,
                                     l_seq=40 if not cigar else None))
            ql = sum(n_ for n_, op in cigar if op in "MIS=X")
            good = not flag & 0x4 and not flag & 0x400 and mapq >= 20
            if good and cigar:                    # no CIGAR, no query length: not in the estimator's counter
                hist[ql] = hist.get(ql, 0) + 1
            if good and not flag & 0x80 and ql:
                want[chrom].append((pos, ql, bool(flag & 16)))
    recs.append(W.bam_record(-1, -1, 0, 4, [], b"unplaced", l_seq=30))      # no reference
    return recs, want, hist


@pytest.mark.parametrize("block_payload,threads,aligned,share_bytes", [
    (65280, 1, False, None), (997, 3, False, None), (64, 2, False, None), (8192, 4, False, None),
    (8192, 4, True, None), (997, 3, True, None), (150, 2, True, None),      # 150: some records are longer than a block
    (8192, 4, False, 20000), (997, 3, False, 5000), (8192, 4, True, 20000), (997, 3, True, 5000),
    (64, 2, False, 100), (150, 2, True, 300), (64, 3, False, 1)])           # shares shorter than a record
def test_synthetic_bam_flags_cigars_and_block_boundaries(tmp_path, monkeypatch, block_payload, threads, aligned,
                                                         share_bytes):
    """Every share of a batch but the first starts at a BGZF block boundary and has to guess where
    its first record begins: at the boundary in an htslib-written file (`aligned`), a few bytes on
    where the writer cuts blocks anywhere (htsjdk).  The guesses are validated in file order;
    where one is wrong or impossible (shares inside one long record) the share is filtered again
    from the true boundary -- the result is the same either way.  PMIO_SHARE_BYTES shrinks the
    shares so that a 300 KB file has many of them."""
    if share_bytes is None:
        monkeypatch.delenv("PMIO_SHARE_BYTES", raising=False)
    else:
        monkeypatch.setenv("PMIO_SHARE_BYTES", str(share_bytes))
    recs, want, hist = synthetic_records()
    path = str(tmp_path / "syn.bam")
    W.write_bam(path, [("c1", 200000), ("c2", 5000), ("c3", 90000)], recs, block_payload, align_records=aligned)
    bam = io_native.BamFile(path, mapq_criteria=20, threads=threads)
    if share_bytes in (20000, 5000):
        assert bam.stats["reparsed_shares"] == 0        # all guesses right, aligned or not
    if share_bytes in (100, 1):
        assert bam.stats["reparsed_shares"] > 0         # the fallback ran, too
    assert bam.references == ("c1", "c2", "c3") and bam.lengths == (200000, 5000, 90000)
    assert sorted(bam.reads) == ["c1", "c3"] and bam.sorted
    for chrom in ("c1", "c3"):
        r = bam.reads[chrom]
        got = list(zip(r["pos"].tolist(), r["readlen"].tolist(), r["is_reverse"].tolist()))
        assert got == want[chrom]
    assert bam.readlen_counter == hist
    assert list(bam.readlen_counter) == list(hist)      # lengths in first-seen order (MODE ties, readlen.pyx:81)
    st = bam.stats
    assert st["n_records"] == 4501 and st["n_with_ref"] == 4500
    assert st["n_unmapped"] == sum(1 for i in range(3000) if i % 11 == 2) + sum(1 for i in range(1500) if i % 11 == 2)
    assert st["n_paired"] == st["n_read2"] + sum(1 for n in (3000, 1500) for i in range(n) if i % 11 == 5)
    assert st["n_accepted"] == len(want["c1"]) + len(want["c3"])


def test_unsorted_bam_is_reported_and_refused(tmp_path):
    recs = [W.bam_record(0, 100, 30, 0, [(36, "M")]), W.bam_record(0, 50, 30, 0, [(36, "M")])]
    path = str(tmp_path / "unsorted.bam")
    W.write_bam(path, [("c1", 1000)], recs)
    bam = io_native.BamFile(path, 1)
    assert not bam.sorted
    calc = CCBitArrayCalculator(50, 36, bam.references, bam.lengths, engine=OracleEngine())
    with pytest.raises(ReadUnsortedError):
        io_native.feed_bam(calc, bam)
    # a chromosome revisited after another one
    recs = [W.bam_record(0, 10, 30, 0, [(36, "M")]), W.bam_record(1, 10, 30, 0, [(36, "M")]),
            W.bam_record(0, 20, 30, 0, [(36, "M")])]
    W.write_bam(path, [("c1", 1000), ("c2", 1000)], recs)
    assert not io_native.BamFile(path, 1).sorted


def test_order_is_checked_per_selected_chromosome_like_the_reference(tmp_path):
    """The reference skips reads of excluded chromosomes BEFORE `_check_pos` (handler/calc.py:143-144,
    mscc.pyx:351-366) and feeds chromosomes in file order: a file that is unsorted only inside an
    excluded chromosome feeds fine; a selected chromosome that goes backwards, or is revisited after
    another SELECTED chromosome, raises; chromosomes are fed in the order they first appear."""
    M = [(36, "M")]

    class Recorder(object):
        def __init__(self):
            self.fed = []

        def feed_packed(self, chrom, **kw):
            self.fed.append((chrom, kw["fwd"].tolist()))

    refs = [("c1", 1000), ("c2", 1000), ("c3", 1000)]
    path = str(tmp_path / "order.bam")
    # c3 first (header order differs from file order), c2 unsorted inside, c1 clean
    recs = [W.bam_record(2, 5, 30, 0, M), W.bam_record(2, 9, 30, 0, M),
            W.bam_record(1, 100, 30, 0, M), W.bam_record(1, 50, 30, 0, M),
            W.bam_record(0, 7, 30, 0, M)]
    W.write_bam(path, refs, recs)
    bam = io_native.BamFile(path, 1)
    assert not bam.sorted and bam.ref_order == ["c3", "c2", "c1"]
    assert bam.ref_sorted == {"c1": True, "c2": False, "c3": True}
    with pytest.raises(ReadUnsortedError):
        io_native.feed_bam(Recorder(), bam)
    rec = Recorder()
    io_native.feed_bam(rec, bam, only={"c1", "c3"})                 # c2 excluded: no error, file order
    assert rec.fed == [("c3", [6, 10]), ("c1", [8])]
    with pytest.raises(ReadUnsortedError):
        bam.packed("c2")
    # c1 ... c2 ... c1: revisited -- unless c2 is excluded, then c1 simply continues
    recs = [W.bam_record(0, 10, 30, 0, M), W.bam_record(1, 10, 30, 0, M), W.bam_record(0, 20, 30, 0, M)]
    W.write_bam(path, refs, recs)
    bam = io_native.BamFile(path, 1)
    assert bam.ref_order == ["c1", "c2", "c1"] and bam.ref_sorted["c1"]
    with pytest.raises(ReadUnsortedError):
        io_native.feed_bam(Recorder(), bam)
    rec = Recorder()
    io_native.feed_bam(rec, bam, only={"c1"})
    assert rec.fed == [("c1", [11, 21])]
    # ... but not if c1 went backwards across the excluded stretch
    recs = [W.bam_record(0, 30, 30, 0, M), W.bam_record(1, 10, 30, 0, M), W.bam_record(0, 20, 30, 0, M)]
    W.write_bam(path, refs, recs)
    bam = io_native.BamFile(path, 1)
    with pytest.raises(ReadUnsortedError):
        io_native.feed_bam(Recorder(), bam, only={"c1"})


def test_mode_estimate_breaks_ties_by_first_occurrence_like_the_reference(tmp_path):
    """core/readlen.pyx:81: `[k for k, v in sorted(c.items(), key=lambda x: x[1])][-1]` -- a stable
    sort by count over a dict in insertion order: among equally frequent lengths the one that
    first occurs LAST in the file wins.  Records without a CIGAR are not counted at all."""
    path = str(tmp_path / "mode.bam")
    lens = [50, 36, 50, 36, 75, 75, 40]                   # 50, 36, 75 twice each; first seen 50, 36, 75
    recs = [W.bam_record(0, 10 * i, 30, 0, [(n, "M")]) for i, n in enumerate(lens)]
    recs.append(W.bam_record(0, 500, 30, 0, [], l_seq=20))   # no CIGAR
    W.write_bam(path, [("c1", 10000)], recs)
    bam = io_native.BamFile(path, 1)
    assert list(bam.readlen_counter.items()) == [(50, 2), (36, 2), (75, 2), (40, 1)]
    assert io_native.estimate_readlen(bam, "MODE", 1) == 75
    assert io_native.estimate_readlen(bam, "MIN", 1) == 36      # (0 before: CIGAR-less records counted as length 0)
    lens = [75, 36, 50, 36, 75, 50]                       # first seen 75, 36, 50 -> 50
    W.write_bam(path, [("c1", 10000)], [W.bam_record(0, 10 * i, 30, 0, [(n, "M")]) for i, n in enumerate(lens)])
    assert io_native.estimate_readlen(io_native.BamFile(path, 1), "MODE", 1) == 50


def test_bam_errors(tmp_path):
    with pytest.raises(IOError):
        io_native.BamFile(str(tmp_path / "missing.bam"))
    p = str(tmp_path / "text.bam")
    open(p, "wb").write(b"@HD\tVN:1.6\n" * 10)
    with pytest.raises(ValueError):
        io_native.BamFile(p)
    # truncated in the middle of a block / of a record; corrupted payload (CRC)
    good = str(tmp_path / "good.bam")
    W.write_bam(good, [("c1", 100000)], [W.bam_record(0, i, 30, 0, [(36, "M")]) for i in range(2000)], 4000)
    raw = open(good, "rb").read()
    for cut in (len(raw) // 2, len(raw) - 40):
        open(p, "wb").write(raw[:cut])
        with pytest.raises(ValueError):
            io_native.BamFile(p)
    blob = bytearray(raw)
    blob[len(raw) // 2] ^= 0xFF
    open(p, "wb").write(bytes(blob))
    with pytest.raises(ValueError):
        io_native.BamFile(p)
    # a file that ends inside a record although every block is intact
    stream_cut = str(tmp_path / "cut.bam")
    recs = [W.bam_record(0, 5, 30, 0, [(36, "M")])]
    W.write_bam(stream_cut, [("c1", 1000)], recs + [recs[0][:20]])
    with pytest.raises(ValueError):
        io_native.BamFile(stream_cut)


# ------------------------------------------------------------------------------------- BigWig
def test_reference_bigwig_fixture_decodes_to_the_bedgraph_arrays():
    fx = helpers.load_fixture()
    bw = io_native.BigWigReader(BIGWIG)
    assert bw.chromsizes == {"chr1": 249250621}
    b, e = bw.fetch_arrays(1.0, "chr1")
    assert np.array_equal(b, fx["map_begins"]) and np.array_equal(e, fx["map_ends"])
    rows = bw.fetch(1.0, "chr1")
    assert rows[0] == (int(b[0]), int(e[0]), 1.0) and len(rows) == 985
    every = bw.fetch_arrays(0.0, "chr1", with_values=True)
    assert every[0].size > b.size and every[2].min() < 1.0          # unfiltered: all intervals
    keep = every[2] >= 1.0
    assert np.array_equal(every[0][keep], b)
    with pytest.raises(KeyError):
        bw.fetch(1.0, "chr2")
    bw.disable_progress_bar()
    bw.close()
    bw.close()
    with pytest.raises(ValueError):
        bw.fetch_arrays(1.0, "chr1")


@pytest.mark.parametrize("compress,two_level", [(True, False), (False, False), (True, True)])
def test_synthetic_bigwig_sections(tmp_path, compress, two_level):
    chroms = [("chrB", 50000), ("chrA", 30000), ("chrEmpty", 1000)]
    sections = [
        {"chrom_id": 0, "type": "bedgraph", "items": [(0, 10, 1.0), (10, 25, 0.5), (40, 100, 1.0)]},
        {"chrom_id": 0, "type": "varstep", "span": 5, "items": [(200, 1.0), (300, 0.25), (400, 2.0)]},
        {"chrom_id": 0, "type": "fixedstep", "start": 1000, "step": 10, "span": 4, "items": [1.0, 0.0, 1.0, 1.0]},
        {"chrom_id": 1, "type": "bedgraph", "items": [(5, 6, 1.0), (29990, 30000, 1.0)]},
        {"chrom_id": 1, "type": "bedgraph", "items": [(30000 - 1, 30000 + 50, 1.0)]},   # overhangs the end
    ]
    path = str(tmp_path / "syn.bw")
    W.write_bigwig(path, chroms, sections, compress=compress, two_level=two_level)
    bw = io_native.BigWigReader(path)
    assert bw.chromsizes == {"chrB": 50000, "chrA": 30000, "chrEmpty": 1000}
    b, e, v = bw.fetch_arrays(1.0, "chrB", with_values=True)
    assert list(zip(b.tolist(), e.tolist())) == [(0, 10), (40, 100), (200, 205), (400, 405),
                                                 (1000, 1004), (1020, 1024), (1030, 1034)]
    assert v.tolist() == [1.0, 1.0, 1.0, 2.0, 1.0, 1.0, 1.0]
    b, e = bw.fetch_arrays(0.0, "chrB")
    assert b.size == 10
    b, e = bw.fetch_arrays(1.0, "chrA")
    assert list(zip(b.tolist(), e.tolist())) == [(5, 6), (29990, 30000), (29999, 30050)]   # as stored
    b, e = bw.fetch_arrays(1.0, "chrEmpty")
    assert b.size == 0 and e.size == 0
    with pytest.raises(KeyError):
        bw.fetch_arrays(1.0, "chrZ")
    bw.close()


def test_bigwig_errors(tmp_path):
    with pytest.raises(IOError):
        io_native.BigWigReader(str(tmp_path / "missing.bw"))
    p = str(tmp_path / "junk.bw")
    open(p, "wb").write(b"\0" * 200)
    with pytest.raises(IOError):
        io_native.BigWigReader(p)
    open(p, "wb").write(struct.pack("<I", 0x888FFC26) + b"\xff" * 100)     # offsets outside the file
    with pytest.raises(IOError):
        io_native.BigWigReader(p)
    with pytest.raises(ValueError):
        io_native.BigWigReader("-")
    # a data block that does not inflate: the same ValueError from the per-chromosome decode
    # and from a worker thread of the prefetch
    p = str(tmp_path / "rotten.bw")
    sections = [{"chrom_id": 0, "type": "bedgraph", "items": [(10 * i + 1000 * k, 10 * i + 1000 * k + 5, 1.0) for i in range(50)]}
                for k in range(40)]
    W.write_bigwig(p, [("c1", 100000)], sections)
    blob = bytearray(open(p, "rb").read())
    data_off = struct.unpack_from("<Q", blob, 16)[0]
    blob[data_off + 8 + 300:data_off + 8 + 340] = b"\xff" * 40        # inside the zlib streams of the first blocks
    open(p, "wb").write(bytes(blob))
    r = io_native.BigWigReader(p)
    with pytest.raises(ValueError):
        r.fetch_arrays(1.0, "c1")
    with pytest.raises(ValueError):
        r.prefetch(1.0, 3)
    r.close()


# ----------------------------------------------------------------------------------- pipeline
def test_run_files_from_bam_and_bigwig_matches_golden_tables(tmp_path):
    """pymasc_b200.pipeline.run_files: files in, the reference's golden curves out; a second
    file over the same track shares it (one device batch) and a header / BigWig length
    mismatch resolves to the longer length (reader/bam.py:217-255)."""
    from pymasc_b200 import pipeline
    gold = helpers.load_golden_tables()
    engine = OracleEngine()
    # second file: the same reads shifted by 7 bp on a header that understates chr1
    fx = helpers.load_fixture()
    recs = [W.bam_record(0, int(p) - 1 + 7, 30, 16 if r else 0, [(int(l), "M")])
            for p, l, r in zip(fx["pos"], fx["readlen"], fx["is_reverse"])]
    other = str(tmp_path / "other.bam")
    W.write_bam(other, [("chr1", 249250000), ("chrM", 16571)], recs)
    results, info = pipeline.run_files([BAM, other], BIGWIG, max_shift=300, mapq_criteria=10,
                                       engine=engine)
    assert len(engine.batches) == 1
    assert [f["read_len"] for f in info["files"]] == [36, 36]
    whole = results[0]
    for got, want in ((whole.chroms["chr1"].cc, gold["cc"]["chr1"]),
                      (whole.mappable_chroms["chr1"].cc, gold["mscc"]["chr1"])):
        want = np.array(want)
        assert np.max(np.abs(np.asarray(got) - want) / np.abs(want)) <= 1e-9
    assert results[1].chroms["chr1"].genomelen == 249250621          # the BigWig's, longer
    assert results[1].genomelen == 249250621 + 16571
    alone, _ = pipeline.run_files(other, BIGWIG, max_shift=300, mapq_criteria=10, engine=OracleEngine())
    helpers.assert_snapshots_equal(helpers.snapshot_whole(results[1]), helpers.snapshot_whole(alone[0]))
    with pytest.raises(ValueError):
        pipeline.run_files(BAM, None, max_shift=20, mapq_criteria=10, engine=OracleEngine())
    assert pipeline.reconcile_chromosome_sizes(["a", "b", "c"], [10, 20, 30], {"a": 12, "b": 15}) == [12, 20, 30]


def test_chromosome_filter_selects_what_the_reference_selects():
    """pipeline.filter_chroms against the reference's utils/calc.py:49-91 on 68 include /
    exclude option lists (tests/golden/chromfilter_cases.json, tests/golden/make_chromfilter_cases.py)."""
    import json
    from pymasc_b200 import pipeline
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "chromfilter_cases.json")) as fh:
        gold = json.load(fh)
    assert len(gold["cases"]) >= 60
    for filters, want in gold["cases"]:
        filters = None if filters is None else [(bool(flag), pats) for flag, pats in filters]
        assert sorted(pipeline.filter_chroms(gold["chroms"], filters)) == want, filters
    assert pipeline.apply_chromfilter(["b", "a", "c"], [2, 1, 3], [(False, ["a"])]) == (["b", "c"], [2, 3])
    with pytest.raises(pipeline.NoTargetChromosomes):
        pipeline.apply_chromfilter(["b", "a"], [2, 1], [(True, ["zzz*"])])


def test_run_files_with_a_chromosome_filter(tmp_path):
    """-i / -e as run_files(chromfilter=...): only the selected chromosomes get results, equal to
    their results in the unfiltered run; reads elsewhere are skipped (handler/calc.py:143-144);
    the read length is still estimated over the whole file."""
    from pymasc_b200 import pipeline
    rng = np.random.default_rng(61)
    refs = [("c1", 60000), ("c2", 40000), ("c3_random", 30000)]
    recs = []
    for ref, (_, ln) in enumerate(refs):
        for p in np.sort(rng.integers(1, ln - 100, 400)):
            recs.append(W.bam_record(ref, int(p), 30, 16 if rng.random() < 0.5 else 0, [(36 if ref else 40, "M")]))
    bam = str(tmp_path / "three.bam")
    W.write_bam(bam, refs, recs)
    sections = [{"chrom_id": ci, "type": "bedgraph", "items": [(s, s + 300, 1.0) for s in range(100, ln - 400, 1000)]}
                for ci, (_, ln) in enumerate(refs)]
    bw = str(tmp_path / "three.bigwig")
    W.write_bigwig(bw, refs, sections)
    everything, info_all = pipeline.run_files(bam, bw, max_shift=120, mapq_criteria=10, engine=OracleEngine())
    picked, info = pipeline.run_files(bam, bw, max_shift=120, mapq_criteria=10, engine=OracleEngine(),
                                      chromfilter=[(False, ["*_random"]), (False, ["c1"])])
    assert info["files"][0]["read_len"] == info_all["files"][0]["read_len"]
    assert sorted(picked[0].chroms) == ["c2"] and sorted(picked[0].mappable_chroms) == ["c2"]
    assert picked[0].genomelen == 40000
    for table in ("chroms", "mappable_chroms"):
        a, b = getattr(picked[0], table)["c2"], getattr(everything[0], table)["c2"]
        assert list(a.ccbins) == list(b.ccbins) and a.forward_sum == b.forward_sum and a.reverse_sum == b.reverse_sum
    with pytest.raises(pipeline.NoTargetChromosomes):
        pipeline.run_files(bam, bw, max_shift=120, mapq_criteria=10, engine=OracleEngine(), chromfilter=[(True, ["chrZ"])])


@pytest.mark.parametrize("spec", helpers.TOY_SPECS, ids=lambda s: s["name"])
def test_packed_feed_from_bam_equals_reference_vectors(tmp_path, spec):
    """The toy cases (exact duplicates, strand collisions, variable query lengths, read-less
    chromosomes) written as BAM, decoded natively, reduced by pack_reference and fed with
    feed_packed: the whole result must equal the unmodified reference's, read-length sums
    included (they depend on the duplicate rules of mscc.pyx:388-393 and :416-418)."""
    toy = helpers.make_toy(spec)
    recs = []
    for ci, chrom in enumerate(toy["references"]):
        r = toy["reads"].get(chrom)
        if r is None:
            continue
        for p, l, rev in zip(r["pos"].tolist(), r["readlen"].tolist(), r["is_reverse"].tolist()):
            recs.append(W.bam_record(ci, p - 1, 30, 16 if rev else 0, [(l, "M")]))
    path = str(tmp_path / "toy.bam")
    W.write_bam(path, list(zip(toy["references"], toy["lengths"])), recs, 3000)
    bam = io_native.BamFile(path, 1)
    feeder = helpers.ListFeeder(toy["tracks"]) if toy["tracks"] is not None else None
    calc = CCBitArrayCalculator(spec["D"], spec["rl"], toy["references"], toy["lengths"], feeder,
                                spec["skip_ncc"], engine=OracleEngine())
    io_native.feed_bam(calc, bam)
    calc.finishup_calculation()
    helpers.assert_snapshots_equal(helpers.snapshot_whole(calc.get_whole_result()),
                                   helpers.load_ref_vectors()["toy_" + spec["name"]])
    # per-read arrays and the packed form agree with a numpy restatement of the rules
    for chrom in bam.reads:
        r, k = bam.reads[chrom], bam.packed(chrom)
        f = r["pos"][~r["is_reverse"]]
        keep = np.ones(f.size, bool)
        keep[1:] = f[1:] != f[:-1]
        assert np.array_equal(k["fwd"], f[keep]) and k["fwd_len_sum"] == int(r["readlen"][~r["is_reverse"]][keep].sum())
        ends = (r["pos"] + r["readlen"] - 1)[r["is_reverse"]]
        u, first = np.unique(ends, return_index=True)
        assert np.array_equal(k["rev"], u) and k["rev_len_sum"] == int(r["readlen"][r["is_reverse"]][first].sum())
        assert (k["first_pos"], k["last_pos"]) == (int(r["pos"][0]), int(r["pos"][-1]))


def test_feed_packed_closes_the_chromosome_to_further_reads():
    calc = CCBitArrayCalculator(50, 20, ["c1", "c2"], [1000, 1000], engine=OracleEngine())
    calc.feed_packed("c1", [5, 9], 40, [60], 20, 5, 41)
    with pytest.raises(ValueError):
        calc.feed_forward_read("c1", 100, 20)
    with pytest.raises(ValueError):
        calc.feed_packed("c1", [200], 20, [], 0, 200, 200)
    calc.feed_forward_read("c2", 7, 20)                  # the next chromosome feeds normally
    with pytest.raises(ReadUnsortedError):
        calc.feed_packed("c1", [300], 20, [], 0, 300, 300)     # c1 is already solved


# ------------------------------------------------------------------- mappable-length statistics
def test_mappability_handler_reproduces_the_reference_json(tmp_path):
    """pymasc_b200.mappability.MappabilityHandler on the reference's BigWig fixture must write
    the reference's own tests/data/hg19_36mer-test_mappability.json (committed inside
    tests/golden/golden_encff000rmb.json): same keys, same 266 values per list."""
    import json
    import shutil
    from pymasc_b200.mappability import MappabilityHandler, BWIOError
    want = helpers.load_golden_tables()["mappability_json"]
    bw = str(tmp_path / "hg19_36mer-test.bigwig")
    shutil.copyfile(BIGWIG, bw)
    assert MappabilityHandler.calc_mappable_len_required_shift_size(36, 300) == 265
    assert MappabilityHandler.calc_mappable_len_required_shift_size(36, 50) == 36
    h = MappabilityHandler(bw, max_shift=300, readlen=36, engine=OracleEngine())
    assert h.max_shift == 265 and h.need_save_stats and not h.is_called
    assert str(h.map_path).endswith("hg19_36mer-test_mappability.json")
    h.calc_mappability()
    assert h.is_called and list(h.chrom2mappable_len) == ["chr1"]
    h.save_mappability_stats()
    got = json.load(open(h.map_path))
    assert got == want
    assert h.get_mappable_len("chr1", 0, 3) == tuple(want["references"]["chr1"][:3])
    assert h.get_mappable_len(shift_from=10, shift_to=12) == want["__whole__"][10:12]
    assert h.get_mappable_len("chrNope") is None
    # a second handler finds the cache and calculates nothing; a longer shift needs an update
    engine = OracleEngine()
    h2 = MappabilityHandler(bw, max_shift=200, readlen=36, engine=engine)
    assert h2.is_called and not h2.need_save_stats and h2.max_shift == 165
    assert h2.mappable_len == want["__whole__"][:166]
    h2.calc_mappability()
    assert engine.batches == []
    h3 = MappabilityHandler(bw, max_shift=400, readlen=36, engine=engine)
    assert h3.need_save_stats and not h3.is_called
    assert h3.get_mappable_len("chr1", force=True)[:266] == tuple(want["references"]["chr1"])
    with pytest.raises(BWIOError):
        MappabilityHandler(str(tmp_path / "missing.bw"), 300, 36)
    # the body of pymasc-precalc (calcmappablelen.py:54-58) in one call, to a path of its own
    from pymasc_b200 import pipeline
    out = str(tmp_path / "stats" / "precalc.json")
    done = pipeline.precalc_mappability(bw, max_shift=300, max_readlen=36, stats_path=out, engine=OracleEngine())
    assert json.load(open(out)) == want and str(done.map_path) == out and not done.need_save_stats


def test_run_files_takes_text_sam_too(tmp_path):
    """pysam reads SAM and BAM alike for the reference; run_files sniffs the gzip magic and
    parses text SAM through io_text.  The same reads as SAM and as BAM give the same result."""
    from pymasc_b200 import pipeline
    fx = helpers.load_fixture()
    n = 400
    refs = [("chr1", 249250621), ("chrM", 16571)]
    sam = ["@HD\tVN:1.6\tSO:coordinate"] + ["@SQ\tSN:%s\tLN:%d" % r for r in refs]
    recs = []
    for i, (p, l, r) in enumerate(zip(fx["pos"][:n], fx["readlen"][:n], fx["is_reverse"][:n])):
        flag = (16 if r else 0) | (0x400 if i % 50 == 7 else 0)
        mapq = 5 if i % 40 == 3 else 30
        sam.append("r%d\t%d\tchr1\t%d\t%d\t%dM\t*\t0\t0\t%s\t%s" % (i, flag, p, mapq, l, "A" * l, "I" * l))
        recs.append(W.bam_record(0, int(p) - 1, mapq, flag, [(int(l), "M")]))
    sam.append("u\t4\t*\t0\t0\t*\t*\t0\t0\tACGT\tIIII")
    sam_path, bam_path = str(tmp_path / "x.sam"), str(tmp_path / "x.bam")
    open(sam_path, "w").write("\n".join(sam) + "\n")
    W.write_bam(bam_path, refs, recs)
    a, ia = pipeline.run_files(sam_path, BIGWIG, max_shift=200, mapq_criteria=10, engine=OracleEngine())
    b, ib = pipeline.run_files(bam_path, BIGWIG, max_shift=200, mapq_criteria=10, engine=OracleEngine())
    assert ia["files"][0]["read_len"] == ib["files"][0]["read_len"] == 36
    assert ia["files"][0]["stats"]["n_accepted"] == ib["files"][0]["stats"]["n_accepted"] < n
    helpers.assert_snapshots_equal(helpers.snapshot_whole(a[0]), helpers.snapshot_whole(b[0]))


def test_read_length_estimators_follow_the_reference_rules():
    """core/readlen.pyx:22-86 restated naively (loops over the sorted counter) against the
    vectorised estimators, on random counters including exact even splits."""
    import random

    def median(c):                      # readlen.pyx:41-69
        num, acc = sum(c.values()), 0
        keys = sorted(c)
        for i, l in enumerate(keys):
            acc += c[l]
            if num % 2:
                if num / 2 <= acc:
                    return l
            elif num / 2 < acc:
                return l
            elif num / 2 == acc:
                return int(round((l + float(keys[i + 1])) / 2))

    rnd = random.Random(7)
    for _ in range(3000):
        c = {rnd.randint(20, 60): rnd.randint(1, 6) for _ in range(rnd.randint(1, 6))}
        assert io_native.ESTFUNCTIONS["MEDIAN"](c) == median(c), c
        assert io_native.ESTFUNCTIONS["MEAN"](c) == int(round(sum(l * f for l, f in c.items()) / float(sum(c.values()))))
        assert (io_native.ESTFUNCTIONS["MIN"](c), io_native.ESTFUNCTIONS["MAX"](c)) == (min(c), max(c))
        assert c[io_native.ESTFUNCTIONS["MODE"](c)] == max(c.values())
    assert io_native.ESTFUNCTIONS["MEDIAN"]({35: 2, 36: 2}) == 36        # 35.5 rounds half to even
    assert io_native.ESTFUNCTIONS["MEDIAN"]({36: 2, 37: 2}) == 36        # 36.5 rounds half to even


def test_hostile_bgzf_sizes_are_format_errors(tmp_path):
    """A block that claims more than 64 KiB of inflated data (BGZF allows at most 65536) must not
    size a buffer from that claim; a header that announces 2^31 references must not reserve them."""
    raw = bytearray(W.bgzf_block(b"BAM\x01" + struct.pack("<ii", 0, 1) + struct.pack("<i", 3) + b"c1\0" + struct.pack("<i", 100)))
    p = str(tmp_path / "isize.bam")
    blob = bytearray(raw)
    blob[-4:] = struct.pack("<I", 0xFFFFFFF0)
    open(p, "wb").write(bytes(blob) + W.BGZF_EOF)
    with pytest.raises(ValueError):
        io_native.BamFile(p)
    p2 = str(tmp_path / "nref.bam")
    open(p2, "wb").write(W.bgzf_block(b"BAM\x01" + struct.pack("<ii", 0, 0x7FFFFFFF)) + W.BGZF_EOF)
    with pytest.raises(ValueError):
        io_native.BamFile(p2)               # truncated header, not an allocation of 2^31 names


# ------------------------------------------------------------------ the library's own inflate
def _raw_deflate(data, level, strategy, wbits=-15):
    import zlib
    co = zlib.compressobj(level, zlib.DEFLATED, wbits, 9, strategy)
    return co.compress(data) + co.flush()


def _inflate_payloads():
    rng = np.random.default_rng(11)
    yield b""
    yield b"a"
    yield b"a" * 65536                                           # one long run: distance 1, length 258
    yield bytes(range(256)) * 200
    yield rng.integers(0, 256, 65280, dtype=np.uint8).tobytes()   # incompressible: stored blocks
    yield b"ACGT" * 9000 + rng.integers(0, 256, 100, dtype=np.uint8).tobytes() + b"T" * 7000
    for n in (1, 7, 8, 9, 257, 258, 259, 4095, 30000, 65536):
        yield rng.integers(0, 4, n, dtype=np.uint8).tobytes()
        skew = np.r_[np.full(16, 0.05), np.full(240, 0.2 / 240)]  # long and short codes: second-level tables
        yield rng.choice(np.arange(256, dtype=np.uint8), n, p=skew).tobytes()
        yield (b"SRR1234567.%d\0" % n) * (n // 12 + 1)


def test_own_inflate_agrees_with_zlib_on_every_block_type():
    """pmb_inflate.h decodes what zlib wrote -- stored, fixed-Huffman and dynamic blocks, short
    windows, levels 0-9 -- byte for byte, and the zlib route of the same entry point agrees."""
    import zlib
    n = 0
    for data in _inflate_payloads():
        for level in (0, 1, 6, 9):
            for strategy in (zlib.Z_DEFAULT_STRATEGY, zlib.Z_FIXED, zlib.Z_HUFFMAN_ONLY, zlib.Z_RLE):
                for wbits in (-15, -9):
                    stream = _raw_deflate(data, level, strategy, wbits)
                    assert io_native.inflate_raw(stream, len(data)) == data
                    n += 1
        assert io_native.inflate_raw(_raw_deflate(data, 6, zlib.Z_DEFAULT_STRATEGY), len(data), use_zlib=True) == data
    assert n > 1000


def test_own_inflate_refuses_what_it_cannot_vouch_for():
    """Wrong size, truncation and flipped bits: the decoder either refuses (the BGZF path then asks
    zlib and the CRC decides) or returns exactly what zlib returns; it never writes past the
    buffer it was given (the canary after it stays intact inside the library call)."""
    import zlib
    rng = np.random.default_rng(12)
    data = rng.integers(0, 16, 30000, dtype=np.uint8).tobytes()
    stream = _raw_deflate(data, 6, zlib.Z_DEFAULT_STRATEGY)
    for wrong in (len(data) - 1, len(data) + 1, 0):
        with pytest.raises(ValueError):
            io_native.inflate_raw(stream, wrong)
    with pytest.raises(ValueError):
        io_native.inflate_raw(b"\x07", 10)                      # reserved block type 3
    refused = 0
    for _ in range(600):
        b = bytearray(stream)
        b[int(rng.integers(len(b)))] ^= 1 << int(rng.integers(8))
        if rng.random() < 0.3:
            b = b[:int(rng.integers(len(b)))]
        try:
            got = io_native.inflate_raw(bytes(b), len(data))
        except ValueError:
            refused += 1
            continue
        try:
            assert got == io_native.inflate_raw(bytes(b), len(data), use_zlib=True)
        except ValueError:
            pass                                                # zlib refuses it: the CRC check is what decides
    assert refused > 100


def test_bam_decodes_the_same_with_and_without_the_own_inflate(tmp_path):
    """A/B of the BGZF path: PMIO_ZLIB_ONLY=1 (read once per process, so in a child) gives the
    same arrays as the default route."""
    import subprocess
    import sys
    recs, _, _ = synthetic_records()
    p = str(tmp_path / "ab.bam")
    W.write_bam(p, [("c1", 200000), ("c2", 5000), ("c3", 90000)], recs, block_payload=4000, align_records=True)
    code = ("import sys, hashlib; sys.path.insert(0, %r); from pymasc_b200 import io_native; "
            "b = io_native.BamFile(%r, mapq_criteria=1); h = hashlib.sha256(); "
            "[h.update(b.reads[c][k].tobytes()) for c in sorted(b.reads) for k in ('pos', 'readlen', 'is_reverse')]; "
            "print(h.hexdigest(), b.stats['n_accepted'])"
            % (os.path.dirname(os.path.dirname(os.path.abspath(__file__))), p))
    outs = []
    for env_extra in ({}, {"PMIO_ZLIB_ONLY": "1"}):
        env = {k: v for k, v in os.environ.items() if k != "PMIO_ZLIB_ONLY"}
        env.update(env_extra)
        outs.append(subprocess.run([sys.executable, "-c", code], env=env, check=True,
                                   capture_output=True, text=True).stdout.strip())
    assert outs[0] == outs[1] and int(outs[0].split()[1]) > 1000


def test_blocks_cut_anywhere_decode_in_parallel_without_a_second_pass(tmp_path):
    """An htsjdk-style file (records straddle the BGZF blocks) of several MiB at the default share
    size: every share finds its first record from the bytes alone, nothing is filtered twice, and
    the arrays equal those of the htslib-style file with the same records."""
    rng = np.random.default_rng(21)
    n = 120000
    pos = np.cumsum(rng.integers(0, 30, n)) + 1
    recs = [W.bam_record(0 if i < n // 2 else 1, int(pos[i] if i < n // 2 else pos[i] - pos[n // 2] + 1) - 1,
                         int(rng.integers(0, 60)), 16 if rng.random() < 0.5 else 0,
                         [(int(rng.integers(30, 51)), "M")], b"SRR%07d.%d" % (i * 7919 % 10**7, i))
            for i in range(n)]
    refs = [("chrA", 4_000_000), ("chrB", 4_000_000)]
    loaded = []
    for name, aligned in (("cut.bam", False), ("whole.bam", True)):
        p = str(tmp_path / name)
        W.write_bam(p, refs, recs, block_payload=65280, align_records=aligned)
        assert os.path.getsize(p) > 2_000_000
        bam = io_native.BamFile(p, mapq_criteria=10, threads=4)
        assert bam.stats["reparsed_shares"] == 0 and bam.stats["n_records"] == n and bam.sorted
        loaded.append(bam)
    for chrom in ("chrA", "chrB"):
        for key in ("pos", "readlen", "is_reverse"):
            assert np.array_equal(loaded[0].reads[chrom][key], loaded[1].reads[chrom][key])
        assert all(np.array_equal(x, y) if isinstance(x, np.ndarray) else x == y
                   for x, y in zip(loaded[0].packed(chrom), loaded[1].packed(chrom)))


def test_clmul_crc32_agrees_with_zlib_on_every_length():
    """pmb_crc32.h (carry-less multiplication, folding by 512 and 128 bits, the last 16..31 bytes
    through zlib) against zlib.crc32: every length around the 16- and 64-byte strides, BGZF block
    sizes, constant and random bytes."""
    import zlib
    rng = np.random.default_rng(31)
    lengths = list(range(0, 330)) + [1000, 4095, 4096, 4097, 65279, 65280, 65536, (1 << 20) + 7]
    for n in lengths:
        d = rng.integers(0, 256, n, dtype=np.uint8).tobytes()
        assert io_native.crc32(d) == zlib.crc32(d) == io_native.crc32(d, use_zlib=True), n
    for n in (64, 65, 127, 128, 4096, 100001):
        for fill in (b"\0", b"\xff", b"\x80"):
            assert io_native.crc32(fill * n) == zlib.crc32(fill * n)
    assert io_native.crc32(b"123456789" * 8) == zlib.crc32(b"123456789" * 8)
    assert io_native.crc32(b"123456789", use_zlib=True) == 0xCBF43926      # the check value of CRC-32


@pytest.mark.parametrize("threads", [1, 3, 0])
def test_bigwig_prefetch_equals_the_per_chromosome_decode(tmp_path, threads):
    """pmio_bw_prefetch decodes the whole file on a thread pool; fetch_arrays / fetch with the
    prefetched threshold return the same arrays as the per-chromosome decode, another threshold
    goes through the decoder again, a chromosome without data comes back empty."""
    rng = np.random.default_rng(41)
    chroms = [("c1", 500000), ("c2", 300000), ("empty", 1000), ("c3", 800000)]
    sections = []
    for ci, (_, ln) in enumerate(chroms):
        if ci == 2:
            continue
        starts = np.sort(rng.choice(ln - 50, 4000, replace=False))
        starts = starts[np.r_[True, np.diff(starts) > 40]]
        for o in range(0, len(starts), 100):            # many blocks: several pieces per chromosome
            items = [(int(s), int(s) + int(rng.integers(1, 40)), float(rng.choice([0.5, 1.0]))) for s in starts[o:o + 100]]
            sections.append({"chrom_id": ci, "type": "bedgraph", "items": items})
    p = str(tmp_path / "many.bigwig")
    W.write_bigwig(p, chroms, sections, two_level=True)
    plain, pre = io_native.BigWigReader(p), io_native.BigWigReader(p)
    pre.prefetch(1.0, threads)
    for name, _ in chroms:
        a, b = plain.fetch_arrays(1.0, name, with_values=True), pre.fetch_arrays(1.0, name, with_values=True)
        assert all(np.array_equal(x, y) for x, y in zip(a, b)) and (len(a[0]) > 0) == (name != "empty")
        assert plain.fetch(1.0, name) == pre.fetch(1.0, name)
        a, b = plain.fetch_arrays(0, name, with_values=True), pre.fetch_arrays(0, name, with_values=True)
        assert all(np.array_equal(x, y) for x, y in zip(a, b)) and (len(a[0]) > len(pre.fetch_arrays(1.0, name)[0]) or name == "empty")
    pre.prefetch(0, threads)                              # replaces the cache
    assert all(np.array_equal(x, y) for x, y in zip(plain.fetch_arrays(0, "c3"), pre.fetch_arrays(0, "c3")))
    with pytest.raises(KeyError):
        pre.fetch_arrays(0, "nope")
    pre.close()
    with pytest.raises(ValueError):
        pre.prefetch(1.0)
    plain.close()


def test_two_stream_inflate_decodes_each_stream_on_its_own_terms():
    """pmb_inflate::inflate2 interleaves the decode loops of two streams (pairs of BGZF blocks):
    every combination of block types and sizes comes out as zlib wrote it, and a corrupt or
    truncated partner does not disturb the other stream."""
    import zlib
    payloads = list(_inflate_payloads())
    rng = np.random.default_rng(51)
    streams = []
    for k, data in enumerate(payloads):
        level = (0, 1, 6, 9)[k % 4]
        strategy = (zlib.Z_DEFAULT_STRATEGY, zlib.Z_FIXED, zlib.Z_HUFFMAN_ONLY, zlib.Z_RLE)[(k // 4) % 4]
        streams.append((_raw_deflate(data, level, strategy), data))
    n = 0
    for i in range(len(streams)):
        for j in rng.choice(len(streams), 6, replace=False):
            (ca, a), (cb, b) = streams[i], streams[int(j)]
            assert io_native.inflate_raw2(ca, len(a), cb, len(b)) == (a, b)
            n += 1
    assert n > 150
    (ca, a), (cb, b) = [(c, d) for c, d in streams if len(d) > 20000 and 2000 < len(c) < len(d) // 2][:2]
    assert io_native.inflate_raw2(ca, len(a), cb[:len(cb) // 2], len(b)) == (a, None)
    assert io_native.inflate_raw2(ca[:100], len(a), cb, len(b)) == (None, b)
    assert io_native.inflate_raw2(ca, len(a) + 1, cb, len(b) - 1) == (None, None)
    refused = 0
    for _ in range(300):
        bad = bytearray(ca)
        bad[int(rng.integers(len(bad)))] ^= 1 << int(rng.integers(8))
        got_a, got_b = io_native.inflate_raw2(bytes(bad), len(a), cb, len(b))
        assert got_b == b
        if got_a is None:
            refused += 1
        else:
            try:
                assert got_a == io_native.inflate_raw(bytes(bad), len(a), use_zlib=True)
            except ValueError:
                pass
    assert refused > 30


def _load_outcome(path, threads, share_bytes, monkeypatch):
    if share_bytes is None:
        monkeypatch.delenv("PMIO_SHARE_BYTES", raising=False)
    else:
        monkeypatch.setenv("PMIO_SHARE_BYTES", str(share_bytes))
    try:
        bam = io_native.BamFile(path, mapq_criteria=20, threads=threads)
    except (ValueError, IOError) as e:
        return ("error", type(e).__name__)
    out = [bam.stats["n_records"], bam.stats["n_accepted"], bam.sorted, sorted(bam.readlen_counter.items())]
    for chrom in sorted(bam.reads):
        r = bam.reads[chrom]
        out.append((chrom, r["pos"].tobytes(), r["readlen"].tobytes(), r["is_reverse"].tobytes()))
    bam.close()
    return out


def test_damaged_bam_files_give_the_same_outcome_however_the_work_is_split(tmp_path, monkeypatch):
    """Bit flips in the compressed file, truncation, and bytes changed in the INFLATED stream
    (re-wrapped in valid BGZF blocks, so only the record parser can notice): the loader either
    refuses the file or returns arrays -- and which of the two, and which arrays, must not depend
    on threads and share size, because every share's guessed start is validated against the
    sequential order of the records."""
    import random
    import zlib
    rnd = random.Random(71)
    recs, _, _ = synthetic_records()
    refs = [("c1", 200000), ("c2", 5000), ("c3", 90000)]
    base, mutated = str(tmp_path / "base.bam"), str(tmp_path / "m.bam")
    refused = 0
    for it in range(36):
        W.write_bam(base, refs, recs, rnd.choice([150, 997, 4000, 65280]), align_records=it % 2 == 0)
        blob = bytearray(open(base, "rb").read())
        mode = it % 3
        if mode == 0:
            for _ in range(rnd.randint(1, 3)):
                blob[rnd.randrange(len(blob))] ^= 1 << rnd.randrange(8)
            data = bytes(blob)
        elif mode == 1:
            data = bytes(blob[:rnd.randrange(1, len(blob))])
        else:
            o, payloads = 0, []
            while o < len(blob):
                bsize = struct.unpack_from("<H", blob, o + 16)[0] + 1
                payloads.append(bytearray(zlib.decompress(bytes(blob[o + 18:o + bsize - 8]), -15)))
                o += bsize
            for _ in range(rnd.randint(1, 4)):
                k = rnd.randrange(len(payloads))
                if payloads[k]:
                    payloads[k][rnd.randrange(len(payloads[k]))] = rnd.randrange(256)
            data = b"".join(W.bgzf_block(bytes(p)) for p in payloads if p) + W.BGZF_EOF
        open(mutated, "wb").write(data)
        sequential = _load_outcome(mutated, 1, 1 << 30, monkeypatch)
        for threads, share in ((3, 3000), (4, 100), (2, None)):
            assert _load_outcome(mutated, threads, share, monkeypatch) == sequential, (it, mode, threads, share)
        refused += sequential[0] == "error"
    assert 5 < refused < 36


def test_bigwig_decodes_the_same_with_and_without_the_own_inflate(tmp_path):
    """BigWig data blocks are zlib streams of unknown inflated size: the library's decoder
    (inflate_upto + Adler-32) and zlib's uncompress (PMIO_ZLIB_ONLY=1, read once per process, so
    in a child) must give the same arrays, on the reference's track and on a file with every
    section type."""
    import subprocess
    import sys
    chroms = [("c1", 50000), ("c2", 30000)]
    sections = [{"chrom_id": 0, "type": "bedgraph", "items": [(100 * i, 100 * i + 37, 1.0 if i % 3 else 0.25) for i in range(400)]},
                {"chrom_id": 1, "type": "varstep", "items": [(50 * i, 1.0) for i in range(300)], "span": 20},
                {"chrom_id": 1, "type": "fixedstep", "items": [1.0, 0.5] * 100, "start": 20000, "step": 30, "span": 10}]
    mixed = str(tmp_path / "mixed.bigwig")
    W.write_bigwig(mixed, chroms, sections)
    code = ("import sys, hashlib; sys.path.insert(0, %r); from pymasc_b200 import io_native; h = hashlib.sha256(); n = 0\n"
            "for path in %r:\n"
            "    for pre in (False, True):\n"
            "        r = io_native.BigWigReader(path)\n"
            "        if pre: r.prefetch(1.0, 2)\n"
            "        for c in sorted(r.chromsizes):\n"
            "            for a in r.fetch_arrays(1.0, c, with_values=True) + r.fetch_arrays(0, c, with_values=True):\n"
            "                h.update(a.tobytes()); n += a.size\n"
            "        r.close()\n"
            "print(h.hexdigest(), n)"
            % (os.path.dirname(os.path.dirname(os.path.abspath(__file__))), [BIGWIG, mixed]))
    outs = []
    for env_extra in ({}, {"PMIO_ZLIB_ONLY": "1"}):
        env = {k: v for k, v in os.environ.items() if k != "PMIO_ZLIB_ONLY"}
        env.update(env_extra)
        outs.append(subprocess.run([sys.executable, "-c", code], env=env, check=True,
                                   capture_output=True, text=True).stdout.strip())
    assert outs[0] == outs[1] and int(outs[0].split()[1]) > 10000
