"""The N > 1 paths on real devices (skipped with fewer than two GPUs in the box; the host logic of
the same paths runs on the CPU over gloo in tests/test_sharding.py): position shards summed by
pmb_allreduce of the C ABI -- one process driving n contexts (pmb_comm_init_all) and one process
per GPU under torchrun (pmb_comm_get_id / pmb_comm_init)."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

import helpers
from oracle import cc_oracle
from pymasc_b200 import _shim
from pymasc_b200.exceptions import PositionOutOfRange

pytestmark = pytest.mark.gpu


def n_devices():
    try:
        return _shim.device_count()
    except Exception:       # noqa: BLE001
        return 0


def case(seed, L, rl, n):
    rng = np.random.default_rng(seed)
    fwd = np.unique(rng.integers(1, L, size=n)).astype(np.uint32)
    rev = np.unique(rng.integers(rl, L + rl - 1, size=n)).astype(np.uint32)
    edges = np.sort(rng.choice(np.arange(0, L), size=2 * max(2, L // 700), replace=False))
    return fwd, rev, edges[0::2].astype(np.uint32), edges[1::2].astype(np.uint32)


@pytest.mark.skipif(n_devices() < 2, reason="needs two GPUs")
def test_one_process_position_shards_are_summed_on_the_devices():
    from pymasc_b200.multigpu import MultiDeviceEngine
    rl, D = 36, 1300
    n = min(n_devices(), 4)
    jobs, wants = [], []
    for c, L in enumerate([350000, 70000, 260000]):
        fwd, rev, b, e = case(800 + c, L, rl, L // 10)
        jobs.append(_shim.Job(L, rl, D, 3, fwd, rev, b, e))
        wants.append(cc_oracle.chrom_counts(L, rl, D, fwd, rev, b, e))
    eng = MultiDeviceEngine(list(range(n)), split="position")
    try:
        for _ in range(2):
            got = eng.run(jobs)
            for g, w in zip(got, wants):
                for k in w:
                    assert np.array_equal(np.asarray(g[k]), np.asarray(w[k])), k
        bad = list(jobs)
        # (a reverse 5' end: the shard owning the chromosome's end keeps reads up to hi + max_shift)
        bad[0] = _shim.Job(350000, rl, D, 3, jobs[0].fwd_pos, np.append(jobs[0].rev_pos, np.uint32(350000 + rl + D + 100)),
                           jobs[0].map_begins, jobs[0].map_ends)
        with pytest.raises(PositionOutOfRange):
            eng.run(bad)
        got = eng.run(jobs)
        for g, w in zip(got, wants):
            for k in w:
                assert np.array_equal(np.asarray(g[k]), np.asarray(w[k])), k
    finally:
        eng.close()


@pytest.mark.skipif(n_devices() < 2, reason="needs two GPUs")
def test_one_process_per_gpu_sharded_engine_under_torchrun():
    n = 2
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n),
           "--master-addr", "127.0.0.1", "--master-port", "29671",
           os.path.join(helpers.ROOT, "tests", "sharded_script.py")]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, (p.stdout[-2000:], p.stderr[-3000:])
    lines = [json.loads(l[7:]) for l in p.stdout.splitlines() if l.startswith("RESULT ")]
    assert sorted(x["rank"] for x in lines) == list(range(n))
    for x in lines:
        assert x["equal_0"] is True and x["equal_1"] is True, x
        assert x["unsorted_equal"] is True, x
        assert x["range_error"] == "raised", x
        assert x["equal_after_error"] is True, x


@pytest.mark.skipif(n_devices() < 2, reason="needs two GPUs")
def test_repeated_execute_and_allreduce_of_resident_batches():
    """What bench.py's resident arm does at N > 1: the SAME batch is executed and all-reduced in
    place again and again.  The reduce leaves the sum in every rank's counter block -- also in the
    vectors of chromosomes outside the rank's range, which the next execute must not build on."""
    from pymasc_b200.distributed import shard_jobs
    rl, D = 36, 500
    n = 2
    jobs, wants = [], []
    for c, L in enumerate([300000, 60000, 200000, 90000]):
        fwd, rev, b, e = case(900 + c, L, rl, L // 10)
        flags = 3 if c != 1 else 1
        jobs.append(_shim.Job(L, rl, D, flags, fwd, rev, b if flags & 2 else None, e if flags & 2 else None))
        wants.append(cc_oracle.chrom_counts(L, rl, D, fwd, rev, b if flags & 2 else None, e if flags & 2 else None))
    devs = [_shim.Device(k) for k in range(n)]
    try:
        _shim.Device.comm_init_all(devs)
        for path in (3, 2):
            batches = []
            for k, d in enumerate(devs):
                d.set_option("path", path)
                batches.append(d.prepare(shard_jobs(jobs, k, n)))
            for it in range(4):
                for b_ in batches:
                    b_.execute()
                _shim.Device.allreduce(batches)
                for b_ in batches:
                    got = b_.fetch()
                    for g, w in zip(got, wants):
                        for key in w:
                            assert np.array_equal(np.asarray(g[key]), np.asarray(w[key])), (key, it, path)
            for b_ in batches:
                b_.free()
    finally:
        for d in devs:
            d.close()
