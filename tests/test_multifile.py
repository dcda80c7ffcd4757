"""Multi-file runs over one mappability track (BASELINE config 5): the merged batch must answer
every calculator exactly as if it had been finished alone.  Host logic here (OracleEngine stands
in for the device); the device side of FLAG_SAME_TRACK is tested in tests/test_gpu_parity.py."""
import numpy as np
import pytest

import helpers
from oracle_engine import OracleEngine
from pymasc_b200 import CCBitArrayCalculator, MultiFileCalculation, _shim
from pymasc_b200.multifile import merge_shared_tracks


def toy_files(spec, n_files):
    """n_files read sets over the tracks of one toy."""
    toys = []
    for f in range(n_files):
        s = dict(spec, seed=spec["seed"] + 100 * f)
        toys.append(helpers.make_toy(s))
    for t in toys[1:]:
        t["tracks"] = toys[0]["tracks"]
    return toys


def feed(calc, toy):
    for chrom, r in toy["reads"].items():
        calc.feed_batch(chrom, r["pos"], r["readlen"], r["is_reverse"])


@pytest.mark.parametrize("spec", [helpers.TOY_SPECS[0], helpers.TOY_SPECS[1], helpers.TOY_SPECS[7]],
                         ids=lambda s: s["name"])
def test_merged_run_equals_separate_runs(spec):
    toys = toy_files(spec, 3)
    # file 2 has no reads on the first chromosome: its job there is an autocorrelation fill
    del toys[2]["reads"]["c1"]
    feeder = helpers.ListFeeder(toys[0]["tracks"])
    engine = OracleEngine()
    alone, together = [], []
    for t in toys:
        for dst, eng in ((alone, OracleEngine()), (together, engine)):
            c = CCBitArrayCalculator(spec["D"], spec["rl"], t["references"], t["lengths"], feeder,
                                     spec["skip_ncc"], None, None, engine=eng)
            feed(c, t)
            dst.append(c)
    for c in alone:
        c.finishup_calculation()
    MultiFileCalculation(together).finishup_calculation()
    assert len(engine.batches) == 1                      # one device batch for all files
    for a, b in zip(alone, together):
        helpers.assert_snapshots_equal(helpers.snapshot_whole(b.get_whole_result()),
                                       helpers.snapshot_whole(a.get_whole_result()))


def test_merge_forms_runs_and_collapses_fills():
    F, M, A, S = _shim.FLAG_NCC, _shim.FLAG_MSCC, _shim.FLAG_AUTOCORR, _shim.FLAG_SAME_TRACK
    b, e = np.array([1, 50], np.uint32), np.array([20, 90], np.uint32)

    def job(flags, L, key):
        j = _shim.Job(L, 10, 30, flags, [3], [40], b if key else None, e if key else None)
        j.track_key = key
        return j

    lists = [
        [job(F | M, 1000, ("t", "c1")), job(F | M, 2000, ("t", "c2")), job(A, 500, ("t", "c3"))],
        [job(F | M, 1000, ("t", "c1")), job(F, 2000, None), job(A, 500, ("t", "c3"))],
        [job(F | M, 1000, ("u", "c1")), job(F | M, 2000, ("t", "c2"))],      # another track on c1
    ]
    merged, route = merge_shared_tracks(lists)
    flags = [(j.chrom_len, j.flags) for j in merged]
    assert flags == [(1000, F | M), (1000, F | M | S), (2000, F | M), (2000, F | M | S),
                     (500, A), (2000, F), (1000, F | M)]
    assert route == [[0, 2, 4], [1, 5, 4], [6, 3]]
    assert merged[1].map_begins.size == 0 and merged[0].map_begins.size == 2
    assert merged[1].fwd_pos is lists[1][0].fwd_pos          # reads are not copied


def test_calculators_with_different_geometry_get_their_own_batches():
    spec = helpers.TOY_SPECS[0]
    toys = toy_files(spec, 2)
    feeder = helpers.ListFeeder(toys[0]["tracks"])
    engine = OracleEngine()
    calcs = []
    for t, rl in zip(toys, (20, 24)):
        c = CCBitArrayCalculator(spec["D"], rl, t["references"], t["lengths"], feeder, False,
                                 engine=engine)
        feed(c, t)
        calcs.append(c)
    MultiFileCalculation(calcs).finishup_calculation()
    assert len(engine.batches) == 2
    for c, t, rl in zip(calcs, toys, (20, 24)):
        ref = CCBitArrayCalculator(spec["D"], rl, t["references"], t["lengths"], feeder, False,
                                   engine=OracleEngine())
        feed(ref, t)
        ref.finishup_calculation()
        helpers.assert_snapshots_equal(helpers.snapshot_whole(c.get_whole_result()),
                                       helpers.snapshot_whole(ref.get_whole_result()))


# ------------------------------------------------------------------ several devices, one process
def _batch_for_devices():
    from pymasc_b200.multifile import merge_shared_tracks
    spec = helpers.TOY_SPECS[0]
    toys = toy_files(spec, 3)
    feeder = helpers.ListFeeder(toys[0]["tracks"])
    calcs = []
    for t in toys:
        c = CCBitArrayCalculator(spec["D"], spec["rl"], t["references"], t["lengths"], feeder, False,
                                 engine=OracleEngine())
        feed(c, t)
        c._close_chromosome()
        calcs.append(c)
    lists = [c._collect_jobs(fill=list(c.references))[0] for c in calcs]
    return merge_shared_tracks(lists)[0]


def test_placement_keeps_shared_track_runs_on_one_device():
    from pymasc_b200.multigpu import place_jobs, job_cost
    jobs = _batch_for_devices()
    S = _shim.FLAG_SAME_TRACK
    for n_dev in (1, 2, 3, 8):
        placement = place_jobs(jobs, n_dev)
        assert sorted(i for idx in placement for i in idx) == list(range(len(jobs)))
        where = {i: d for d, idx in enumerate(placement) for i in idx}
        for i, j in enumerate(jobs):
            if j.flags & S:
                assert where[i] == where[i - 1]             # a follower sits right behind its owner
                assert placement[where[i]].index(i) == placement[where[i]].index(i - 1) + 1
    two = place_jobs(jobs, 2)
    loads = [sum(job_cost(jobs[i]) for i in idx) for idx in two]
    assert min(loads) > 0.5 * max(loads)                    # LPT: no device is left nearly idle


def test_multi_device_engine_equals_one_device():
    from pymasc_b200.multigpu import MultiDeviceEngine
    jobs = _batch_for_devices()
    want = OracleEngine().run(jobs)
    devices = [OracleEngine(), OracleEngine(), OracleEngine()]
    eng = MultiDeviceEngine(devices)
    got = eng.run(jobs)
    assert sum(len(d.batches) for d in devices) == 3 and all(d.batches for d in devices)
    for g, w in zip(got, want):
        assert sorted(g) == sorted(w)
        for k in w:
            assert np.array_equal(np.asarray(g[k]), np.asarray(w[k])), k
    assert eng.run([]) == []

    class Broken(object):
        def run(self, jobs):
            raise RuntimeError("device lost")
    with pytest.raises(RuntimeError):
        MultiDeviceEngine([OracleEngine(), Broken()]).run(jobs)


def test_multifile_calculation_over_several_devices():
    from pymasc_b200.multigpu import MultiDeviceEngine
    spec = helpers.TOY_SPECS[0]
    toys = toy_files(spec, 3)
    feeder = helpers.ListFeeder(toys[0]["tracks"])
    alone, together = [], []
    for t in toys:
        for dst in (alone, together):
            c = CCBitArrayCalculator(spec["D"], spec["rl"], t["references"], t["lengths"], feeder,
                                     spec["skip_ncc"], engine=OracleEngine())
            feed(c, t)
            dst.append(c)
    for c in alone:
        c.finishup_calculation()
    MultiFileCalculation(together, engine=MultiDeviceEngine([OracleEngine(), OracleEngine()])).finishup_calculation()
    for a, b in zip(alone, together):
        helpers.assert_snapshots_equal(helpers.snapshot_whole(b.get_whole_result()),
                                       helpers.snapshot_whole(a.get_whole_result()))


def test_position_shards_of_a_shared_track_batch_add_up():
    """ShardedEngine's decomposition cuts every job at its own place, so followers are turned
    back into ordinary jobs over the owner's intervals; the rank partials still sum to the whole."""
    from pymasc_b200 import sharding
    from pymasc_b200.distributed import shard_jobs
    jobs = _batch_for_devices()
    assert any(j.flags & _shim.FLAG_SAME_TRACK for j in jobs)
    whole = OracleEngine().run(jobs)
    for world in (2, 3):
        parts = []
        for r in range(world):
            mine = shard_jobs(jobs, r, world)
            assert not any(j.flags & _shim.FLAG_SAME_TRACK for j in mine)
            parts.append(OracleEngine().run(mine))
        for a, m in zip(whole, sharding.sum_outputs(parts)):
            assert sorted(a) == sorted(m)
            for k in a:
                assert np.array_equal(np.asarray(a[k]), np.asarray(m[k])), k


def test_multi_device_engine_position_split_equals_one_device():
    from pymasc_b200.multigpu import MultiDeviceEngine
    jobs = _batch_for_devices()
    want = OracleEngine().run(jobs)
    for n_dev in (2, 3):
        devices = [OracleEngine() for _ in range(n_dev)]
        got = MultiDeviceEngine(devices, split="position").run(jobs)
        assert all(d.batches == [len(jobs)] for d in devices)        # every device sees every job
        for g, w in zip(got, want):
            assert sorted(g) == sorted(w)
            for k in w:
                assert np.array_equal(np.asarray(g[k]), np.asarray(w[k])), k
    with pytest.raises(ValueError):
        MultiDeviceEngine([OracleEngine()], split="files")


def test_readers_of_the_same_bigwig_file_share_a_track():
    """The reference opens one BigWigReader per input file; two reader objects of one file must
    still form a shared-track run (keyed by path), readers of different files must not."""
    import os
    from pymasc_b200 import io_native
    bw = os.path.join(helpers.GOLDEN, "hg19_36mer-test.bigwig")
    fx = helpers.load_fixture()
    engine = OracleEngine()
    seen = []
    engine_run = engine.run
    engine.run = lambda jobs: (seen.append([j.flags for j in jobs]), engine_run(jobs))[1]
    calcs = []
    for f in range(2):
        c = CCBitArrayCalculator(100, 36, ["chr1"], [fx["lengths"][0]], io_native.BigWigReader(bw), engine=engine)
        c.feed_batch("chr1", fx["pos"] + 3 * f, fx["readlen"], fx["is_reverse"])
        calcs.append(c)
    MultiFileCalculation(calcs).finishup_calculation()
    assert seen == [[3, 3 | _shim.FLAG_SAME_TRACK]]
    assert calcs[0]._track_key("chr1") == calcs[1]._track_key("chr1")
    other = CCBitArrayCalculator(100, 36, ["chr1"], [1000], helpers.ListFeeder({}), engine=engine)
    assert other._track_key("chr1") != calcs[0]._track_key("chr1")
