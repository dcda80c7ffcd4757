"""One process per GPU: position-sharded counts + one all-reduce (torch.distributed).

``ShardedEngine`` is an engine for ``CCBitArrayCalculator`` (``engine=`` keyword).  Every rank
of the process group feeds the same reads (or at least closes the same chromosomes) and calls
``run(jobs)`` with the same job list; each rank counts its position shard of every chromosome
(``sharding.plan_weighted``), the (max_shift+1)-long int64 count vectors are summed across ranks with a
single all-reduce -- ``pmb_allreduce`` of the C ABI (NCCL over NVLink/NVSwitch) when the local
engine is a device, on the device counter block itself (no host round trip) -- and every rank
returns the full result.

torch.distributed is used for the rendezvous only (rank / world, carrying the communicator id,
agreeing that no rank failed before the collective); counts and their sum come from
libpymasc_b200.so.
"""
import numpy as np

from . import _shim, sharding

_KEYS = ("ncc", "dmp", "mfwd", "mrev", "mscc")


class _DevicePointer(object):
    """__cuda_array_interface__ view of a device int64 block owned by a pmb_batch."""

    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {
            "shape": (int(n),), "typestr": "<i8", "data": (int(ptr), False), "version": 3,
            "strides": None,
        }


def counters_tensor(batch):
    """torch int64 CUDA tensor aliasing the batch's device counter block
    (int64[n_jobs][5][stride] followed by int64[n_jobs][2] strand totals)."""
    import torch
    ptr, n, _ = batch.counters()
    return torch.as_tensor(_DevicePointer(ptr, n), device="cuda:%d" % batch.dev.ordinal)


def job_chroms(jobs):
    """The sharding module's chromosome dicts from whole-chromosome jobs.  Position shards cut
    every job at its own place, so a FLAG_SAME_TRACK follower (whose range must equal its
    owner's, include/pymasc_b200.h) becomes an ordinary job over the owner's intervals."""
    out, track = [], (None, None)
    for j in jobs:
        has_map = bool(j.flags & (_shim.FLAG_MSCC | _shim.FLAG_AUTOCORR))
        if has_map and not j.flags & _shim.FLAG_SAME_TRACK:
            track = (j.map_begins, j.map_ends)
        out.append({"length": j.chrom_len, "fwd": j.fwd_pos, "rev": j.rev_pos,
                    "begins": track[0] if has_map else None, "ends": track[1] if has_map else None,
                    "flags": j.flags & ~_shim.FLAG_SAME_TRACK})
    return out


def shard_jobs(jobs, rank, world):
    """This rank's position shards of `jobs` (same length and order as `jobs`)."""
    if not jobs:
        return []
    rl, D = jobs[0].read_len, jobs[0].max_shift
    chroms = job_chroms(jobs)
    pieces = sharding.plan_weighted(chroms, rl, D, world)[rank]     # every rank holds the same jobs
    mine = sharding.jobs_for_rank(chroms, pieces, rl, D, flags_of=lambda c: c["flags"])
    for j in mine:   # read-less autocorrelation jobs carry no reads
        if j.flags & _shim.FLAG_AUTOCORR:
            j.fwd_pos = j.rev_pos = np.zeros(0, np.uint32)
    return mine


class ShardedEngine(object):
    """Engine that splits every batch across the ranks of a torch.distributed group.

    local: a ``_shim.Device`` (product: NCCL all-reduce of the device counters) or any object
    with ``run(jobs)`` returning host dicts (tests: gloo all-reduce of host vectors)."""

    def __init__(self, local=None, group=None):
        import torch.distributed as dist
        self._dist = dist
        self._group = group
        self._local = local
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)

    @property
    def local(self):
        if self._local is None:
            self._local = _shim.Device()      # LOCAL_RANK / PYMASC_B200_DEVICE
        return self._local

    def run(self, jobs):
        if not jobs:
            return []
        mine = shard_jobs(jobs, self.rank, self.world)
        if isinstance(self.local, _shim.Device):
            return self._run_device(mine)
        return self._run_host(mine)

    def _agree(self, failed, device=None):
        """Has ANY rank failed so far?  (A rank that raises before the collective would leave the
        others waiting in it: everyone learns of the failure here and raises instead.)"""
        import torch
        where = ("cuda:%d" % device) if device is not None and self._dist.get_backend(self._group) == "nccl" else "cpu"
        t = torch.tensor([1 if failed else 0], dtype=torch.int32, device=where)
        self._dist.all_reduce(t, op=self._dist.ReduceOp.MAX, group=self._group)
        return bool(int(t.item()))

    def _ensure_comm(self, dev):
        """The context's NCCL communicator (pmb_comm_init): rank 0 of the group creates the id,
        torch.distributed carries its 128 bytes to the others -- the only use of torch here."""
        if getattr(dev, "comm_world", 1) == self.world and getattr(dev, "_comm_group", None) is self._group:
            return
        import torch
        backend = self._dist.get_backend(self._group)
        where = ("cuda:%d" % dev.ordinal) if backend == "nccl" else "cpu"
        t = torch.zeros(_shim.Device.COMM_ID_BYTES, dtype=torch.uint8, device=where)
        if self.rank == 0:
            t.copy_(torch.frombuffer(bytearray(_shim.Device.comm_get_id()), dtype=torch.uint8))
        src = self._dist.get_global_rank(self._group, 0) if self._group is not None else 0
        self._dist.broadcast(t, src=src, group=self._group)
        dev.comm_init(bytes(t.cpu().numpy().tobytes()), self.rank, self.world)
        dev._comm_group = self._group

    # NCCL path: the device counter blocks are summed in place through the C ABI (pmb_allreduce),
    # then one D2H.  A position out of range on any rank raises PositionOutOfRange on every rank.
    def _run_device(self, mine):
        dev = self.local
        self._ensure_comm(dev)
        batch, err = None, None
        try:
            batch = dev.prepare(mine)
            batch.execute()
        except Exception as e:      # noqa: BLE001 -- raised below, once every rank knows
            err = e
        try:
            if self._agree(err is not None, dev.ordinal):
                raise err if err is not None else _shim.DeviceError(
                    "another rank failed before the all-reduce; this rank's counts are discarded")
            batch.allreduce()
            return batch.fetch()
        finally:
            if batch is not None:
                batch.free()

    # host path (gloo): used by the CPU tests of the sharding logic
    def _run_host(self, mine):
        import torch
        outs, err = None, None
        try:
            outs = self.local.run(mine)
        except Exception as e:      # noqa: BLE001
            err = e
        if self._agree(err is not None):
            raise err if err is not None else RuntimeError("another rank failed before the all-reduce")
        flat, index = [], []
        for i, o in enumerate(outs):
            for k in sorted(o):
                v = np.atleast_1d(np.asarray(o[k], dtype=np.int64))
                index.append((i, k, v.size, np.isscalar(o[k]) or isinstance(o[k], int)))
                flat.append(v)
        t = torch.from_numpy(np.concatenate(flat))
        self._dist.all_reduce(t, group=self._group)
        red, off, res = t.numpy(), 0, [dict() for _ in outs]
        for i, k, n, scalar in index:
            res[i][k] = int(red[off]) if scalar else red[off:off + n].copy()
            off += n
        return res
