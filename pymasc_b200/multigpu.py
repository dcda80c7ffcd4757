"""One process, several B200s, whole chromosomes per GPU -- the reference's ``-p N`` unit.

The reference parallelises by chromosome: N worker processes pull chromosomes from a queue
(PyMaSC/handler/calc.py:163-192, PyMaSC/utils/calc.py:119-130) and nothing is reduced across
them, because results are per chromosome.  ``MultiDeviceEngine`` is the same decomposition for
one host process driving N GPUs (BASELINE configs[2] "sharded by chromosome", configs[4] "whole
box"): the jobs of a batch are packed onto the devices longest-processing-time-first, each
device runs its sub-batch through its own ``pmb_ctx`` on its own host thread (ctypes drops the
GIL; contexts of different devices are independent, include/pymasc_b200.h), and the outputs are
put back in job order.  No collective, no device-to-device traffic.

A run of ``FLAG_SAME_TRACK`` jobs (several read sets over one mappability track,
pymasc_b200.multifile) is the unit that is placed -- it stays on one device so that the track is
built once.

``split="position"`` is the other decomposition of SURVEY.md 8(e), for deep libraries whose
largest chromosome is too much for one GPU's share (BASELINE configs[3]): every device counts
one position range of the concatenated genome with halos rebuilt from the clipped inputs
(pymasc_b200.sharding), and the per-device counter blocks -- (max_shift+1) int64 per counter and
chromosome, ~1 MB per device -- are summed in place by ``pmb_allreduce`` (one NCCL communicator
over the process's contexts, ``pmb_comm_init_all``); device 0's block is then the only D2H.
(Several contexts on ONE device -- the single-GPU tests -- cannot form a communicator and are
summed on the host.  One process per GPU is ``pymasc_b200.distributed.ShardedEngine``.)
"""
import os
import threading

from . import _shim
from .sharding import array_bits

# relative cost of one read / one interval against one bit of array, as in sharding.plan_weighted
READ_COST = 150
EDGE_COST = 150


def job_cost(job):
    return (array_bits(job.chrom_len, job.read_len, job.max_shift) if not job.hi else job.hi - job.lo) + \
        READ_COST * (job.fwd_pos.size + job.rev_pos.size) + EDGE_COST * 2 * job.map_begins.size


def place_jobs(jobs, n_devices):
    """LPT packing of the batch's units (a job, or a job with its FLAG_SAME_TRACK followers) onto
    `n_devices` bins.  Returns one sorted list of job indices per device."""
    units = []
    for i, j in enumerate(jobs):
        if j.flags & _shim.FLAG_SAME_TRACK and units:
            units[-1].append(i)
        else:
            units.append([i])
    load = [0] * n_devices
    out = [[] for _ in range(n_devices)]
    for unit in sorted(units, key=lambda u: -sum(job_cost(jobs[i]) for i in u)):
        d = min(range(n_devices), key=lambda k: load[k])
        load[d] += sum(job_cost(jobs[i]) for i in unit)
        out[d].extend(unit)
    for idx in out:
        idx.sort()
    return out


class MultiDeviceEngine(object):
    """Engine for ``CCBitArrayCalculator`` / ``MultiFileCalculation`` (``engine=`` keyword).

    devices: device ordinals (default: every visible CUDA device, or the comma-separated
    ``PYMASC_B200_DEVICES``), or objects with ``run(jobs)`` (tests)."""

    def __init__(self, devices=None, split="chromosome"):
        if split not in ("chromosome", "position"):
            raise ValueError("split must be 'chromosome' or 'position'")
        self.split = split
        if devices is None:
            env = os.environ.get("PYMASC_B200_DEVICES")
            devices = [int(x) for x in env.split(",")] if env else list(range(_shim.device_count()))
            if not devices:
                raise _shim.DeviceError("no CUDA device visible -- pymasc_b200 has no CPU fallback")
        self._specs = list(devices)
        self._engines = [None] * len(self._specs)
        self._comm_ready = False
        self.last_placement = None
        self.last_device_ms = None      # per device: CUDA-event time of its execute (None for test engines)

    def _engine(self, k):
        if self._engines[k] is None:
            spec = self._specs[k]
            self._engines[k] = spec if hasattr(spec, "run") else _DeviceRunner(int(spec))
        return self._engines[k]

    def run(self, jobs):
        if not jobs:
            return []
        n_dev = len(self._specs)
        by_position = self.split == "position" and n_dev > 1
        if by_position:
            from .distributed import shard_jobs
            shards = [shard_jobs(jobs, k, n_dev) for k in range(n_dev)]
            placement = self.last_placement = [list(range(len(jobs)))] * n_dev
            parts = [None] * n_dev
        else:
            placement = self.last_placement = place_jobs(jobs, n_dev)
        results = [None] * len(jobs)
        errors = []
        self.last_device_ms = [None] * n_dev

        # position ranges on distinct devices: counters stay on the devices and are all-reduced
        on_device = by_position and all(not hasattr(sp, "run") for sp in self._specs) and \
            len(set(int(sp) for sp in self._specs)) == n_dev
        batches = [None] * n_dev

        def work(k):
            idx = placement[k]
            if not idx:
                return
            try:
                eng = self._engine(k)
                if on_device:
                    batches[k] = eng.device.prepare(shards[k])
                    batches[k].execute()
                elif by_position:
                    parts[k] = eng.run(shards[k])
                else:
                    for i, out in zip(idx, eng.run([jobs[i] for i in idx])):
                        results[i] = out
                if isinstance(eng, _DeviceRunner):
                    self.last_device_ms[k] = eng.device.timing()["total_ms"]
            except BaseException as e:      # noqa: BLE001 -- re-raised on the calling thread
                errors.append(e)

        threads = [threading.Thread(target=work, args=(k,)) for k in range(1, len(self._specs))]
        for t in threads:
            t.start()
        work(0)
        for t in threads:
            t.join()
        try:
            if errors:
                raise errors[0]
            if on_device:
                if not self._comm_ready:
                    _shim.Device.comm_init_all([self._engine(k).device for k in range(n_dev)])
                    self._comm_ready = True
                _shim.Device.allreduce(batches)
                return batches[0].fetch()
        finally:
            for b in batches:
                if b is not None:
                    b.free()
        if by_position:
            from .sharding import sum_outputs
            return sum_outputs(parts)
        return results

    def close(self):
        for e in self._engines:
            if isinstance(e, _DeviceRunner):
                e.close()
        self._engines = [None] * len(self._specs)
        self._comm_ready = False


class _DeviceRunner(object):
    def __init__(self, ordinal):
        self.device = _shim.Device(ordinal)

    def run(self, jobs):
        return self.device.run_batch(jobs)

    def close(self):
        self.device.close()
