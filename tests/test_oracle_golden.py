"""Pin the CPU oracle (oracle/cc_oracle.c, oracle/np_oracle.py) to the reference.

Pins, in order of authority:
  1. the reference's own golden tables (tests/golden/ENCFF000RMB-test_{cc,mscc,nreads,stats}.tab,
     tests/data/hg19_36mer-test_mappability.json), committed verbatim in
     tests/golden/golden_encff000rmb.json by tests/golden/make_golden.py;
  2. outputs of the compiled, unmodified reference calculator on the fixture and on toy
     cases (tests/golden/ref_vectors.json, same script);
  3. live: the compiled reference in oracle/_ref, when it is present.
None of these tests needs a GPU.
"""
import numpy as np
import pytest

import helpers
from oracle import cc_oracle, np_oracle
from pymasc_b200.result import MSCCResult, NCCResult

D, RL = 300, 36   # the golden command: pymasc ... -d 300 -q 10 -r 36 (test_golden_outputs.py:48-57)


@pytest.fixture(scope="module")
def fx():
    return helpers.load_fixture()


@pytest.fixture(scope="module")
def gold():
    return helpers.load_golden_tables()


@pytest.fixture(scope="module")
def counts(fx):
    chr1_len = fx["lengths"][fx["references"].index("chr1")]
    rev = fx["is_reverse"]
    fwd = fx["pos"][~rev]
    rend = fx["pos"][rev] + fx["readlen"][rev] - 1
    return chr1_len, cc_oracle.chrom_counts(chr1_len, RL, D, fwd, rend, fx["map_begins"], fx["map_ends"])


def test_fixture_shape(fx):
    assert len(fx["references"]) == 86 and sum(fx["lengths"]) == 3137454505
    assert fx["pos"].size == 1292 and fx["map_begins"].size == 985


def test_totals_match_golden_stats(counts, gold):
    _, c = counts
    assert c["fwd_total"] == int(gold["stats"]["Forward reads"]) == 622
    assert c["rev_total"] == int(gold["stats"]["Reverse reads"]) == 670
    assert gold["nreads"]["chr1"][0] == "622-670"


def test_mscc_read_counts_match_golden_nreads(counts, gold):
    _, c = counts
    rows = gold["nreads"]["chr1"][1:]          # row 0 is "raw"
    assert len(rows) == D + 1
    want_f = [int(r.split("-")[0]) for r in rows]
    want_r = [int(r.split("-")[1]) for r in rows]
    assert c["mfwd"].tolist() == want_f
    assert c["mrev"].tolist() == want_r


def test_mappable_len_matches_reference_json(counts, gold):
    _, c = counts
    ml = cc_oracle.mappable_len_list(c["dmp"], RL, D)
    assert len(ml) == D - RL + 2
    want = gold["mappability_json"]["references"]["chr1"]      # max_shift 265 -> 266 entries
    assert ml[:len(want)] == want[:len(ml)]
    assert ml[RL - 1] == int(gold["stats"]["DMP length"]) == 19906
    assert c["mfwd"][RL - 1 - (RL - 1)] >= 0
    # "reads in DMP" are the per-shift sums at shift read_len... the stats row is at index rl-1
    assert int(gold["stats"]["Forward reads in DMP"]) == 385
    assert int(gold["stats"]["Reverse reads in DMP"]) == 441


def test_ncc_curve_matches_golden_cc_tab(counts, gold):
    L, c = counts
    r = NCCResult(max_shift=D, read_len=RL, genomelen=L, forward_sum=c["fwd_total"],
                  reverse_sum=c["rev_total"], forward_read_len_sum=0, reverse_read_len_sum=0,
                  ccbins=c["ncc"].tolist())
    r.calc_cc()
    want = np.array(gold["cc"]["chr1"])
    assert np.max(np.abs(r.cc - want) / np.abs(want)) <= 1e-9       # north_star tolerance
    assert np.array_equal(r.cc, want)                               # in fact bit-identical


def test_mscc_curve_matches_golden_mscc_tab(counts, gold):
    L, c = counts
    r = MSCCResult(max_shift=D, read_len=RL, genomelen=L, forward_sum=c["mfwd"].tolist(),
                   reverse_sum=c["mrev"].tolist(), forward_read_len_sum=0, reverse_read_len_sum=0,
                   ccbins=c["mscc"].tolist(), mappable_len=cc_oracle.mappable_len_list(c["dmp"], RL, D))
    r.calc_cc()
    want = np.array(gold["mscc"]["chr1"])
    assert np.max(np.abs(r.cc - want) / np.abs(want)) <= 1e-9
    assert np.array_equal(r.cc, want)


def test_counts_equal_compiled_reference_vectors(counts):
    _, c = counts
    vec = helpers.load_ref_vectors()["fixture_both"]
    assert c["ncc"].tolist() == vec["chroms"]["chr1"]["ccbins"]
    m = vec["mappable_chroms"]["chr1"]
    assert c["mscc"].tolist() == m["ccbins"]
    assert c["mfwd"].tolist() == m["forward_sum"]
    assert c["mrev"].tolist() == m["reverse_sum"]
    assert cc_oracle.mappable_len_list(c["dmp"], RL, D) == m["mappable_len"]


def _toy_counts(spec, chrom, mod):
    toy = helpers.make_toy(spec)
    r = toy["reads"][chrom]
    rev = r["is_reverse"]
    L = toy["lengths"][toy["references"].index(chrom)]
    trk = toy["tracks"].get(chrom) if toy["tracks"] else None
    return mod.chrom_counts(L, spec["rl"], spec["D"], r["pos"][~rev],
                            r["pos"][rev] + r["readlen"][rev] - 1,
                            trk[0] if trk else None, trk[1] if trk else None)


@pytest.mark.parametrize("spec", helpers.TOY_SPECS, ids=lambda s: s["name"])
def test_toy_counts_equal_reference_vectors(spec):
    """C oracle == numpy closed form == compiled reference, on every toy chromosome with reads."""
    vec = helpers.load_ref_vectors()["toy_" + spec["name"]]
    toy = helpers.make_toy(spec)
    for chrom in toy["reads"]:
        c = _toy_counts(spec, chrom, cc_oracle)
        n = _toy_counts(spec, chrom, np_oracle)
        for k in c:
            assert np.array_equal(np.asarray(c[k]), np.asarray(n[k])), (chrom, k)
        if not spec["skip_ncc"]:
            assert c["ncc"].tolist() == vec["chroms"][chrom]["ccbins"]
            assert c["fwd_total"] == vec["chroms"][chrom]["forward_sum"]
            assert c["rev_total"] == vec["chroms"][chrom]["reverse_sum"]
        if "dmp" in c:
            m = vec["mappable_chroms"][chrom]
            assert c["mscc"].tolist() == m["ccbins"]
            assert c["mfwd"].tolist() == m["forward_sum"]
            assert c["mrev"].tolist() == m["reverse_sum"]
            assert cc_oracle.mappable_len_list(c["dmp"], spec["rl"], spec["D"]) == m["mappable_len"]


def test_readless_chromosome_autocorrelation():
    spec = helpers.TOY_SPECS[0]
    toy = helpers.make_toy(spec)
    vec = helpers.load_ref_vectors()["toy_" + spec["name"]]
    b, e = toy["tracks"]["c2"]
    got = cc_oracle.mappable_autocorr(toy["lengths"][1], spec["rl"], spec["D"], b, e)
    assert got.tolist() == vec["mappable_chroms"]["c2"]["mappable_len"]


def test_position_shards_add_up():
    """Sums over j split at arbitrary bit positions add to the whole."""
    spec = helpers.TOY_SPECS[0]
    toy = helpers.make_toy(spec)
    r = toy["reads"]["c1"]
    rev = r["is_reverse"]
    args = (toy["lengths"][0], spec["rl"], spec["D"], r["pos"][~rev],
            r["pos"][rev] + r["readlen"][rev] - 1) + toy["tracks"]["c1"]
    whole = cc_oracle.chrom_counts(*args)
    nbits = cc_oracle.array_bits(toy["lengths"][0], spec["rl"], spec["D"])
    cuts = [0, 777, 2048, 3999, nbits]
    parts = [cc_oracle.chrom_counts(*args, lo=a, hi=b) for a, b in zip(cuts[:-1], cuts[1:])]
    for k in whole:
        assert np.array_equal(np.asarray(whole[k]), sum(np.asarray(p[k]) for p in parts)), k


def test_out_of_range_position_is_an_error_not_an_abort():
    with pytest.raises(IndexError):
        cc_oracle.chrom_counts(1000, 20, 50, [5], [1000 + 20 + 50 + 100])


def test_live_compiled_reference_if_present(fx):
    from oracle import ref_runner
    if not ref_runner.available():
        pytest.skip("oracle/_ref not built here")
    spec = helpers.TOY_SPECS[5]
    toy = helpers.make_toy(spec)
    by = {c: (r["pos"], r["readlen"], r["is_reverse"]) for c, r in toy["reads"].items()}
    calc = ref_runner.run_reference(spec["D"], spec["rl"], toy["references"], toy["lengths"], by,
                                    toy["tracks"], spec["skip_ncc"])
    snap = helpers.snapshot_whole(calc.get_whole_result())
    helpers.assert_snapshots_equal(snap, helpers.load_ref_vectors()["toy_" + spec["name"]])
