"""Regenerate the committed parity fixtures from the reference checkout.

Run in the CPU container (needs /root/reference and oracle/_ref):

    python oracle/build_ref.py && python tests/golden/make_golden.py

Writes, beside this script:
  fixture_encff000rmb.npz   packed inputs of the reference's fixture: chr1 reads of
                            tests/data/ENCFF000RMB-test.sam that pass the reference's read
                            filter at -q 10, and the >= 1.0 intervals of
                            tests/data/hg19_36mer-test.bedGraph; @SQ names and lengths.
  golden_encff000rmb.json   the reference's own golden tables, verbatim numbers:
                            tests/golden/ENCFF000RMB-test_{cc,mscc,nreads,stats}.tab and
                            tests/data/hg19_36mer-test_mappability.json.
  ref_vectors.json          outputs of the compiled, unmodified reference calculator
                            (oracle/_ref) on the fixture (-d 300 -r 36, NCC+MSCC) and on the
                            toy cases of tests/helpers.py (host semantics: duplicates,
                            read-less chromosomes, missing tracks, skip_ncc, list lengths).
  ENCFF000RMB-test.bam, hg19_36mer-test.bigwig
                            the reference's binary test DATA files (tests/data), byte for byte:
                            the BAM / BigWig twins of the .sam / .bedGraph the arrays above were
                            parsed from -- inputs of tests/test_io_native.py (pymasc_b200.io_native
                            must decode them to exactly fixture_encff000rmb.npz).
"""
import json
import os
import shutil
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from pymasc_b200.io_text import read_sam, read_bedgraph, read_chrom_sizes  # noqa: E402
from oracle import ref_runner  # noqa: E402
import helpers  # noqa: E402

REF = os.environ.get("PYMASC_REFERENCE", "/root/reference")
DATA = os.path.join(REF, "tests", "data")
GOLD = os.path.join(REF, "tests", "golden")


def table(path, cols):
    with open(path) as fh:
        rows = [l.rstrip("\n").split("\t") for l in fh]
    idx = [rows[0].index(c) for c in cols]
    return {c: [r[i] for r in rows[1:]] for c, i in zip(cols, idx)}


def copy_binary_fixtures():
    for name in ("ENCFF000RMB-test.bam", "hg19_36mer-test.bigwig"):
        shutil.copyfile(os.path.join(DATA, name), os.path.join(HERE, name))
        os.chmod(os.path.join(HERE, name), 0o644)


def main():
    copy_binary_fixtures()
    refs, lens, reads = read_sam(os.path.join(DATA, "ENCFF000RMB-test.sam"), mapq_criteria=10)
    assert list(reads) == ["chr1"]
    tracks = read_bedgraph(os.path.join(DATA, "hg19_36mer-test.bedGraph"))
    # (the fixture BigWig only carries chr1, whose size equals the BAM header's, so the
    # size reconciliation of handler/calc.py:108-115 changes nothing: genome length 3137454505)
    sizes = read_chrom_sizes(os.path.join(DATA, "hg19.chrom.sizes"))
    assert sizes["chr1"] == lens[refs.index("chr1")]
    r1 = reads["chr1"]
    np.savez_compressed(
        os.path.join(HERE, "fixture_encff000rmb.npz"),
        references=np.array(refs), lengths=np.array(lens, np.int64),
        chr1_pos=r1["pos"].astype(np.int32), chr1_readlen=r1["readlen"].astype(np.int16),
        chr1_is_reverse=r1["is_reverse"],
        chr1_map_begins=tracks["chr1"][0], chr1_map_ends=tracks["chr1"][1])

    gold = {}
    for name in ("cc", "mscc"):
        t = table(os.path.join(GOLD, "ENCFF000RMB-test_%s.tab" % name), ["shift", "whole", "chr1"])
        gold[name] = {"shift": [int(x) for x in t["shift"]],
                      "whole": [float(x) for x in t["whole"]],
                      "chr1": [float(x) for x in t["chr1"]]}
    t = table(os.path.join(GOLD, "ENCFF000RMB-test_nreads.tab"), ["shift", "whole", "chr1"])
    gold["nreads"] = t
    with open(os.path.join(GOLD, "ENCFF000RMB-test_stats.tab")) as fh:
        gold["stats"] = dict(l.rstrip("\n").split("\t") for l in fh)
    with open(os.path.join(DATA, "hg19_36mer-test_mappability.json")) as fh:
        gold["mappability_json"] = json.load(fh)
    with open(os.path.join(HERE, "golden_encff000rmb.json"), "w") as fh:
        json.dump(gold, fh, indent=0)

    vec = {}
    by_chrom = {"chr1": (r1["pos"], r1["readlen"], r1["is_reverse"])}
    for label, trk, skip in (("fixture_ncc", None, False),
                             ("fixture_both", {"chr1": tracks["chr1"]}, False),
                             ("fixture_mscc_only", {"chr1": tracks["chr1"]}, True)):
        calc = ref_runner.run_reference(300, 36, refs, lens, by_chrom, trk, skip)
        snap = helpers.snapshot_whole(calc.get_whole_result())
        # keep the file small: only chr1 and one read-less contig of the 86 references
        for k in ("chroms", "mappable_chroms"):
            if k in snap:
                snap[k] = {c: v for c, v in snap[k].items() if c in ("chr1", "chr2", "chrM")}
        vec[label] = snap
    for spec in helpers.TOY_SPECS:
        toy = helpers.make_toy(spec)
        by_chrom = {c: (r["pos"], r["readlen"], r["is_reverse"]) for c, r in toy["reads"].items()}
        calc = ref_runner.run_reference(spec["D"], spec["rl"], toy["references"], toy["lengths"],
                                        by_chrom, toy["tracks"], spec["skip_ncc"])
        vec["toy_" + spec["name"]] = helpers.snapshot_whole(calc.get_whole_result())
        # the -p N worker protocol: flush(chrom) + get_result(chrom) per chromosome
        # (handler/worker.py:224-234), one fresh calculator fed a single chromosome
        per = {}
        for c in toy["references"]:
            sub = {c: by_chrom[c]} if c in by_chrom else {}
            wc = ref_runner.run_reference(spec["D"], spec["rl"], toy["references"], toy["lengths"],
                                          sub, toy["tracks"], spec["skip_ncc"], finish=False)
            wc.flush(c)
            res = wc.get_result(c)
            per[c] = {"chrom": helpers.snapshot_chrom(res.chrom),
                      "mappable_chrom": helpers.snapshot_chrom(res.mappable_chrom)}
        vec["toy_" + spec["name"] + "_worker"] = per
    with open(os.path.join(HERE, "ref_vectors.json"), "w") as fh:
        json.dump(vec, fh)
    print("wrote fixtures:", {k: os.path.getsize(os.path.join(HERE, k)) for k in os.listdir(HERE)})


if __name__ == "__main__":
    main()
