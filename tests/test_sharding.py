"""Position-range sharding (the N > 1 path) on the CPU: the plan covers every bit once, the
per-shard partial counts add up bit-exactly to the whole-chromosome counts, and the
world_size-2 ``ShardedEngine`` (gloo) returns the full result on every rank."""
import os
import socket
import sys

import numpy as np
import pytest

import helpers  # noqa: F401  (sys.path)
from oracle import cc_oracle
from oracle_engine import OracleEngine
from pymasc_b200 import _shim, sharding


def toy_chroms(seed=3, lengths=(9000, 700, 12345, 4096), rl=20, with_map=True):
    rng = np.random.default_rng(seed)
    out = []
    for L in lengths:
        fwd = np.unique(rng.integers(1, L, size=L // 15)).astype(np.uint32)
        rev = np.unique(rng.integers(rl, L + rl - 1, size=L // 15)).astype(np.uint32)
        b = e = None
        if with_map:
            edges = np.sort(rng.choice(np.arange(0, L), size=2 * max(2, L // 300), replace=False))
            b, e = edges[0::2].astype(np.uint32), edges[1::2].astype(np.uint32)
        out.append({"length": L, "fwd": fwd, "rev": rev, "begins": b, "ends": e})
    return out


@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_plan_covers_every_bit_once(world):
    lengths, rl, D = [5000, 130, 77777, 1024], 36, 300
    plan = sharding.plan(lengths, rl, D, world)
    assert len(plan) == world
    for ci, L in enumerate(lengths):
        nb = sharding.array_bits(L, rl, D)
        segs = sorted((lo, hi) for pieces in plan for c, lo, hi in pieces if c == ci)
        assert segs[0][0] == 0 and segs[-1][1] == nb
        for (a, b), (c, d) in zip(segs, segs[1:]):
            assert b == c
    sizes = [sum(hi - lo for _, lo, hi in pieces) for pieces in plan]
    assert max(sizes) - min(sizes) <= 2 * sharding.ALIGN_BITS


@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_weighted_plan_covers_every_bit_once_and_balances_events(world):
    """Cuts that equalise bits + 150 * events: all reads of a skewed genome sit in 10 % of it."""
    rng = np.random.default_rng(5)
    rl, D = 36, 300
    chroms = []
    for L in (400000, 90000, 250000):
        hot = np.sort(rng.integers(1, L // 10, size=L // 20)).astype(np.uint32)
        chroms.append({"length": L, "fwd": hot, "rev": hot + 40, "begins": None, "ends": None})
    plan = sharding.plan_weighted(chroms, rl, D, world)
    assert len(plan) == world
    cost = []
    for pieces in plan:
        c = 0.0
        for ci, lo, hi in pieces:
            a = chroms[ci]["fwd"]
            n = int(np.searchsorted(a, hi) - np.searchsorted(a, lo))
            a = chroms[ci]["rev"]
            n += int(np.searchsorted(a, hi) - np.searchsorted(a, lo))
            c += (hi - lo) + sharding.READ_COST * n
        cost.append(c)
    for ci, ch in enumerate(chroms):
        nb = sharding.array_bits(ch["length"], rl, D)
        segs = sorted((lo, hi) for pieces in plan for c, lo, hi in pieces if c == ci)
        assert segs[0][0] == 0 and segs[-1][1] == nb
        for (a, b), (c, d) in zip(segs, segs[1:]):
            assert b == c
    # balanced to within the reads of one 128-bit granule (dense pile-ups: ~20 reads per granule here)
    assert max(cost) - min(cost) <= 0.02 * sum(cost) / world + 2 * sharding.ALIGN_BITS * (1 + sharding.READ_COST)
    equal = sharding.plan([c["length"] for c in chroms], rl, D, world)
    if world > 1:
        assert plan != equal


def test_plan_by_chromosome_is_lpt():
    lengths = [100, 90, 50, 40, 30, 10]
    plan = sharding.plan_by_chromosome(lengths, 10, 20, 2)
    assert sorted(ci for pieces in plan for ci, _, _ in pieces) == list(range(6))
    loads = [sum(hi - lo for _, lo, hi in pieces) for pieces in plan]
    assert abs(loads[0] - loads[1]) <= 130 + 40


@pytest.mark.parametrize("world", [2, 5])
@pytest.mark.parametrize("rl,D", [(20, 100), (36, 30)])
def test_shard_partials_sum_to_whole(world, rl, D):
    chroms = toy_chroms(rl=rl)
    eng = OracleEngine()
    whole = eng.run(sharding.jobs_for_rank(chroms, sharding.plan([c["length"] for c in chroms], rl, D, 1)[0], rl, D))
    parts = []
    for r in range(world):
        pieces = sharding.plan([c["length"] for c in chroms], rl, D, world)[r]
        jobs = sharding.jobs_for_rank(chroms, pieces, rl, D)
        assert len(jobs) == len(chroms)
        parts.append(eng.run(jobs))
    merged = sharding.sum_outputs(parts)
    for a, b in zip(whole, merged):
        assert sorted(a) == sorted(b)
        for k in a:
            assert np.array_equal(np.asarray(a[k]), np.asarray(b[k])), k


def test_sharded_autocorrelation_sums_to_whole():
    rl, D, L = 20, 150, 20000
    c = toy_chroms(lengths=(L,), rl=rl)[0]
    want = cc_oracle.mappable_autocorr(L, rl, D, c["begins"], c["ends"])
    nb = sharding.array_bits(L, rl, D)
    cuts = [0, 4096, 12800, nb]
    got = sum(cc_oracle.mappable_autocorr(L, rl, D, c["begins"], c["ends"], lo=a, hi=b)
              for a, b in zip(cuts, cuts[1:]))
    assert np.array_equal(want, got)


# ---------------------------------------------------------------- world_size 2 over gloo
def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank),
                      WORLD_SIZE=str(world))
    here = os.path.dirname(os.path.abspath(__file__))
    for p in (os.path.dirname(here), here):
        if p not in sys.path:
            sys.path.insert(0, p)
    import torch.distributed as dist
    from oracle_engine import OracleEngine
    from pymasc_b200 import CCBitArrayCalculator
    from pymasc_b200.distributed import ShardedEngine
    import helpers
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        fx = helpers.load_fixture()
        feeder = helpers.ListFeeder({"chr1": (fx["map_begins"], fx["map_ends"])})
        calc = CCBitArrayCalculator(300, 36, fx["references"][:3], fx["lengths"][:3], feeder,
                                    engine=ShardedEngine(local=OracleEngine()))
        calc.feed_batch("chr1", fx["pos"], fx["readlen"], fx["is_reverse"])
        calc.finishup_calculation()
        r = calc.get_result("chr1")
        q.put((rank, r.chrom.ccbins, r.mappable_chrom.ccbins, r.mappable_chrom.forward_sum,
               list(r.mappable_chrom.mappable_len), r.chrom.forward_sum, r.chrom.reverse_sum))
    finally:
        dist.destroy_process_group()


def test_sharded_engine_world2_gloo_matches_golden():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = sorted(q.get(timeout=180) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    vec = helpers.load_ref_vectors()
    assert got[0][1:] == got[1][1:]                     # every rank holds the full result
    _, ncc, mscc, mfwd, ml, ftot, rtot = got[0]
    ref = vec["fixture_both"]                           # the unmodified reference's own output
    assert ncc == ref["chroms"]["chr1"]["ccbins"]
    assert mscc == ref["mappable_chroms"]["chr1"]["ccbins"]
    assert mfwd == ref["mappable_chroms"]["chr1"]["forward_sum"]
    assert ml == ref["mappable_chroms"]["chr1"]["mappable_len"]
    assert (ftot, rtot) == (622, 670)                   # tests/golden/ENCFF000RMB-test_stats.tab
