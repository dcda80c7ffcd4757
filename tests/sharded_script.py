"""Run under torchrun by tests/test_multi_gpu.py: one process per GPU, ShardedEngine behind the
calculator's engine interface -- position shards, pmb_allreduce of the C ABI -- against the oracle.
Every rank prints one JSON line."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "tests"), ROOT):
    sys.path.insert(0, p)

import torch                                           # noqa: E402
import torch.distributed as dist                       # noqa: E402
from oracle import cc_oracle                           # noqa: E402
from pymasc_b200 import _shim                          # noqa: E402
from pymasc_b200.distributed import ShardedEngine      # noqa: E402
from pymasc_b200.exceptions import PositionOutOfRange  # noqa: E402

local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
rank, world = dist.get_rank(), dist.get_world_size()
eng = ShardedEngine(_shim.Device(local))


def case(seed, L, rl, n):
    rng = np.random.default_rng(seed)
    fwd = np.unique(rng.integers(1, L, size=n)).astype(np.uint32)
    rev = np.unique(rng.integers(rl, L + rl - 1, size=n)).astype(np.uint32)
    edges = np.sort(rng.choice(np.arange(0, L), size=2 * max(2, L // 700), replace=False))
    return fwd, rev, edges[0::2].astype(np.uint32), edges[1::2].astype(np.uint32)


rl, D = 36, 700
out = {"rank": rank, "world": world}
jobs, wants = [], []
for c, L in enumerate([400000, 90000, 260000]):
    fwd, rev, b, e = case(40 + c, L, rl, L // 10)
    jobs.append(_shim.Job(L, rl, D, 3, fwd, rev, b, e))
    wants.append(cc_oracle.chrom_counts(L, rl, D, fwd, rev, b, e))
for rep in range(2):                                    # the communicator is reused
    got = eng.run(jobs)
    out["equal_%d" % rep] = all(np.array_equal(np.asarray(g[k]), np.asarray(w[k])) for g, w in zip(got, wants) for k in w)
# unsorted reads in ONE chromosome of this rank's own shard (the host-side clipping of shard_jobs
# presumes sorted lists, so the order is broken after it): the rank finds out on the device while it
# counts, pmb_allreduce's first reduce tells every rank, all recompute and reduce again
from pymasc_b200.distributed import shard_jobs         # noqa: E402
mine = shard_jobs(jobs, rank, world)
if rank == world - 1 and mine[2].fwd_pos.size > 2:
    perm = np.random.default_rng(3).permutation(mine[2].fwd_pos.size)
    mine[2].fwd_pos = np.ascontiguousarray(mine[2].fwd_pos[perm])
try:
    got = eng._run_device(mine)
    out["unsorted_equal"] = all(np.array_equal(np.asarray(g[k]), np.asarray(w[k])) for g, w in zip(got, wants) for k in w)
except Exception as e:      # noqa: BLE001
    out["unsorted_equal"] = "raised %r" % (e,)
# a read beyond the bit array, inside the LAST rank's range: every rank must raise, none may hang
bad = list(jobs)
bad[2] = _shim.Job(260000, rl, D, 3, jobs[2].fwd_pos, np.append(jobs[2].rev_pos, np.uint32(260000 + rl + D + 100)),
                   jobs[2].map_begins, jobs[2].map_ends)
try:
    eng.run(bad)
    out["range_error"] = "no exception"
except PositionOutOfRange:
    out["range_error"] = "raised"
# and the engine still works afterwards
got = eng.run(jobs)
out["equal_after_error"] = all(np.array_equal(np.asarray(g[k]), np.asarray(w[k])) for g, w in zip(got, wants) for k in w)
print("RESULT " + json.dumps(out), flush=True)
dist.barrier()
dist.destroy_process_group()
