"""ctypes binding of libpymasc_b200.so (include/pymasc_b200.h).

The library is the only compute path of this package.  If it is missing, or no sm_100
device can be opened, the calls here raise -- there is deliberately no CPU fallback.
"""
import ctypes
import os
import struct

import numpy as np

from .exceptions import DeviceError, PositionOutOfRange

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PYMASC_B200_LIB") or os.path.join(_HERE, "libpymasc_b200.so")   # override: tuning builds

PMB_OK, PMB_ERR_CUDA, PMB_ERR_RANGE, PMB_ERR_ARG, PMB_ERR_NOMEM = 0, -1, -2, -3, -4
FLAG_NCC, FLAG_MSCC, FLAG_AUTOCORR, FLAG_SAME_TRACK = 1, 2, 4, 8
NCOUNTERS = 5
COUNTER_ORDER = ("ncc", "dmp", "mfwd", "mrev", "mscc")

_u32p = ctypes.POINTER(ctypes.c_uint32)
_i64p = ctypes.POINTER(ctypes.c_int64)


class ChromArgs(ctypes.Structure):
    _fields_ = [
        ("chrom_len", ctypes.c_uint64),
        ("read_len", ctypes.c_uint32),
        ("max_shift", ctypes.c_uint32),
        ("flags", ctypes.c_uint32),
        ("reserved", ctypes.c_uint32),
        ("fwd_pos", ctypes.c_void_p), ("n_fwd", ctypes.c_uint64),
        ("rev_pos", ctypes.c_void_p), ("n_rev", ctypes.c_uint64),
        ("map_begin", ctypes.c_void_p), ("map_end", ctypes.c_void_p),
        ("n_intervals", ctypes.c_uint64),
        ("range_lo", ctypes.c_uint64), ("range_hi", ctypes.c_uint64),
    ]


class ChromOut(ctypes.Structure):
    _fields_ = [
        ("fwd_total", ctypes.c_int64), ("rev_total", ctypes.c_int64),
        ("ncc", ctypes.c_void_p), ("mscc", ctypes.c_void_p), ("mfwd", ctypes.c_void_p),
        ("mrev", ctypes.c_void_p), ("dmp", ctypes.c_void_p),
    ]


class Timing(ctypes.Structure):
    _fields_ = [
        ("build_ms", ctypes.c_float), ("window_ms", ctypes.c_float), ("total_ms", ctypes.c_float),
        ("launches", ctypes.c_uint32), ("window_launches", ctypes.c_uint32),
        ("window_word_shift_pairs", ctypes.c_uint64),
        ("path", ctypes.c_uint32), ("ev_fwd_ms", ctypes.c_float), ("ev_rev_ms", ctypes.c_float),
        ("ev_edge_ms", ctypes.c_float),
    ]


EXPORTS = (
    "pmb_create", "pmb_destroy", "pmb_last_error", "pmb_run_chrom", "pmb_run_batch",
    "pmb_batch_prepare", "pmb_batch_execute", "pmb_batch_fetch", "pmb_batch_counters",
    "pmb_batch_free", "pmb_get_timing", "pmb_stream", "pmb_synchronize", "pmb_host_alloc",
    "pmb_host_free", "pmb_set_option", "pmb_pipe_probe", "pmb_set_stream", "pmb_device_count",
    "pmb_batch_validate", "pmb_run_chrom_noalloc", "pmb_comm_get_id", "pmb_comm_init", "pmb_comm_init_all",
    "pmb_allreduce", "pmb_comm_destroy",
)

_lib = None


def load_library():
    """dlopen the shim (needs libcudart, not a GPU).  Raises DeviceError if it was not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise DeviceError(
            "libpymasc_b200.so is missing (%s). Build it with `python -m pymasc_b200.build`; "
            "pymasc_b200 has no CPU fallback." % LIB_PATH)
    L = ctypes.CDLL(LIB_PATH)
    vp, vpp = ctypes.c_void_p, ctypes.POINTER(ctypes.c_void_p)
    L.pmb_create.argtypes = [vpp, ctypes.c_int]
    L.pmb_create.restype = ctypes.c_int
    L.pmb_device_count.argtypes = []
    L.pmb_device_count.restype = ctypes.c_int
    L.pmb_destroy.argtypes = [vp]
    L.pmb_destroy.restype = None
    L.pmb_last_error.argtypes = [vp]
    L.pmb_last_error.restype = ctypes.c_char_p
    L.pmb_run_chrom.argtypes = [vp, ctypes.POINTER(ChromArgs), ctypes.POINTER(ChromOut)]
    L.pmb_run_chrom.restype = ctypes.c_int
    L.pmb_run_batch.argtypes = [vp, ctypes.POINTER(ChromArgs), ctypes.POINTER(ChromOut), ctypes.c_uint32]
    L.pmb_run_batch.restype = ctypes.c_int
    L.pmb_batch_prepare.argtypes = [vp, ctypes.POINTER(ChromArgs), ctypes.c_uint32, vpp]
    L.pmb_batch_prepare.restype = ctypes.c_int
    L.pmb_batch_execute.argtypes = [vp, vp]
    L.pmb_batch_execute.restype = ctypes.c_int
    L.pmb_batch_fetch.argtypes = [vp, vp, ctypes.POINTER(ChromOut)]
    L.pmb_batch_fetch.restype = ctypes.c_int
    L.pmb_batch_counters.argtypes = [vp, vp, vpp, ctypes.POINTER(ctypes.c_uint64),
                                     ctypes.POINTER(ctypes.c_uint32)]
    L.pmb_batch_counters.restype = ctypes.c_int
    L.pmb_batch_validate.argtypes = [vp, vp]
    L.pmb_batch_validate.restype = ctypes.c_int
    L.pmb_comm_get_id.argtypes = [vp]
    L.pmb_comm_get_id.restype = ctypes.c_int
    L.pmb_comm_init.argtypes = [vp, vp, ctypes.c_int, ctypes.c_int]
    L.pmb_comm_init.restype = ctypes.c_int
    L.pmb_comm_init_all.argtypes = [vpp, ctypes.c_int]
    L.pmb_comm_init_all.restype = ctypes.c_int
    L.pmb_allreduce.argtypes = [vpp, vpp, ctypes.c_int]
    L.pmb_allreduce.restype = ctypes.c_int
    L.pmb_comm_destroy.argtypes = [vp]
    L.pmb_comm_destroy.restype = None
    L.pmb_run_chrom_noalloc.argtypes = [vp, ctypes.POINTER(ChromArgs), vp, ctypes.c_uint64]
    L.pmb_run_chrom_noalloc.restype = ctypes.c_int
    L.pmb_batch_free.argtypes = [vp, vp]
    L.pmb_batch_free.restype = None
    L.pmb_get_timing.argtypes = [vp, ctypes.POINTER(Timing)]
    L.pmb_get_timing.restype = ctypes.c_int
    L.pmb_stream.argtypes = [vp]
    L.pmb_stream.restype = vp
    L.pmb_synchronize.argtypes = [vp]
    L.pmb_synchronize.restype = ctypes.c_int
    L.pmb_set_stream.argtypes = [vp, vp]
    L.pmb_set_stream.restype = ctypes.c_int
    L.pmb_host_alloc.argtypes = [vpp, ctypes.c_uint64]
    L.pmb_host_alloc.restype = ctypes.c_int
    L.pmb_host_free.argtypes = [vp]
    L.pmb_host_free.restype = None
    L.pmb_set_option.argtypes = [vp, ctypes.c_char_p, ctypes.c_int64]
    L.pmb_set_option.restype = ctypes.c_int
    L.pmb_pipe_probe.argtypes = [vp, ctypes.c_int, ctypes.POINTER(ctypes.c_double),
                                 ctypes.POINTER(ctypes.c_float)]
    L.pmb_pipe_probe.restype = ctypes.c_int
    _lib = L
    return L


def _as_u32(a):
    if a is None:
        return np.zeros(0, dtype=np.uint32)
    a = np.asarray(a)
    if a.dtype != np.uint32 or not a.flags.c_contiguous:
        a = np.ascontiguousarray(a, dtype=np.uint32)
    return a


def device_count():
    """CUDA devices the library can see (0 without a driver or GPU)."""
    return int(load_library().pmb_device_count())


class Job(object):
    """One chromosome (or position shard) handed to the device.  ``track_key`` (host only)
    names the mappability track the intervals come from: jobs of one chromosome with equal
    keys may be merged into a FLAG_SAME_TRACK run (pymasc_b200.multifile)."""

    __slots__ = ("chrom_len", "read_len", "max_shift", "flags", "fwd_pos", "rev_pos",
                 "map_begins", "map_ends", "lo", "hi", "track_key", "_rec")

    _FMT = struct.Struct("<QIIIIQQQQQQQQQ")            # pmb_chrom_args, include/pymasc_b200.h

    def __init__(self, chrom_len, read_len, max_shift, flags, fwd_pos=None, rev_pos=None,
                 map_begins=None, map_ends=None, lo=0, hi=0):
        self.chrom_len, self.read_len, self.max_shift, self.flags = (
            int(chrom_len), int(read_len), int(max_shift), int(flags))
        self.fwd_pos, self.rev_pos = _as_u32(fwd_pos), _as_u32(rev_pos)
        self.map_begins, self.map_ends = _as_u32(map_begins), _as_u32(map_ends)
        self.lo, self.hi = int(lo), int(hi)
        self.track_key = None

    def __setattr__(self, name, value):
        object.__setattr__(self, name, value)
        if name != "_rec":
            object.__setattr__(self, "_rec", None)      # the packed record below is stale

    def record(self):
        """This job as the 96 bytes of a pmb_chrom_args (cached until a field is assigned: a
        caller that runs the same jobs again -- resident arrays, another step -- pays nothing).
        The arrays are borrowed: the job keeps them alive."""
        rec = self._rec
        if rec is None:
            f, r, b, e = self.fwd_pos, self.rev_pos, self.map_begins, self.map_ends
            ptr = lambda a: a.__array_interface__["data"][0] if a.size else 0     # noqa: E731
            rec = self._FMT.pack(self.chrom_len, self.read_len, self.max_shift, self.flags, 0, ptr(f), f.size,
                                 ptr(r), r.size, ptr(b), ptr(e), b.size, self.lo, self.hi)
            object.__setattr__(self, "_rec", rec)
        return rec


class Device(object):
    """A pmb_ctx: one CUDA device, one stream."""

    def __init__(self, ordinal=None):
        self._L = load_library()
        if ordinal is None:
            ordinal = int(os.environ.get("PYMASC_B200_DEVICE", os.environ.get("LOCAL_RANK", "0")))
        h = ctypes.c_void_p()
        rc = self._L.pmb_create(ctypes.byref(h), int(ordinal))
        if rc != PMB_OK:
            raise DeviceError("pmb_create(device %d) failed (%d): %s -- pymasc_b200 needs an sm_100 "
                              "GPU and has no CPU fallback" %
                              (ordinal, rc, self._L.pmb_last_error(None).decode()))
        self._h = h
        self.ordinal = int(ordinal)

    def close(self):
        if getattr(self, "_h", None):
            self._L.pmb_destroy(self._h)
            self._h = None

    __del__ = close

    def _check(self, rc):
        if rc == PMB_OK:
            return
        msg = self._L.pmb_last_error(self._h).decode()
        if rc == PMB_ERR_RANGE:
            raise PositionOutOfRange(msg)
        if rc == PMB_ERR_ARG:
            raise ValueError(msg)
        raise DeviceError("libpymasc_b200 error %d: %s" % (rc, msg))

    def set_option(self, name, value):
        self._check(self._L.pmb_set_option(self._h, name.encode(), int(value)))

    def synchronize(self):
        self._check(self._L.pmb_synchronize(self._h))

    @property
    def stream(self):
        return self._L.pmb_stream(self._h)

    def set_stream(self, stream):
        """Adopt a caller-owned cudaStream_t (integer handle; None/0 = the context's own)."""
        self._check(self._L.pmb_set_stream(self._h, ctypes.c_void_p(int(stream) if stream else None)))

    # ---- argument marshalling
    # (one struct.pack_into per job: a dozen ctypes field assignments and four `.ctypes.data`
    # look-ups per job were a third of the host time of a 24-chromosome step)
    _OUT_FMT = struct.Struct("<qqQQQQQ")              # pmb_chrom_out: totals, then ncc, mscc, mfwd, mrev, dmp
    _OUT_ORDER = ("ncc", "mscc", "mfwd", "mrev", "dmp")

    @staticmethod
    def _marshal(jobs):
        n = len(jobs)
        buf = bytearray(b"".join([j.record() for j in jobs]))
        assert len(buf) == n * ctypes.sizeof(ChromArgs)
        args = (ChromArgs * n).from_buffer(buf)
        args._keep = buf
        return args

    @staticmethod
    def _outputs(jobs):
        """Output structs over ONE int64 block (a view per counter): per-array allocations and
        .ctypes lookups cost more host time than the D2H copy itself."""
        n = len(jobs)
        names = []
        total = 0
        for j in jobs:
            if j.flags & FLAG_AUTOCORR:
                ks = ("dmp",)
            else:
                ks = (("ncc",) if j.flags & FLAG_NCC else ()) + \
                     (("dmp", "mfwd", "mrev", "mscc") if j.flags & FLAG_MSCC else ())
            names.append(ks)
            total += len(ks) * (j.max_shift + 1)
        block = np.empty(max(total, 1), np.int64)
        base = block.__array_interface__["data"][0]
        size = Device._OUT_FMT.size
        assert size == ctypes.sizeof(ChromOut)
        buf = bytearray(n * size)
        pack = Device._OUT_FMT.pack_into
        arrays = []
        off = 0
        for i, (j, ks) in enumerate(zip(jobs, names)):
            S = j.max_shift + 1
            d, p = {}, {}
            for k in ks:
                d[k] = block[off:off + S]
                p[k] = base + 8 * off
                off += S
            pack(buf, i * size, 0, 0, *[p.get(k, 0) for k in Device._OUT_ORDER])
            arrays.append(d)
        outs = (ChromOut * n).from_buffer(buf)
        outs._keep = (buf, block)
        return outs, arrays

    @staticmethod
    def _collect(jobs, outs, arrays):
        res = []
        for o, d, j in zip(outs, arrays, jobs):
            d = dict(d)
            if not j.flags & FLAG_AUTOCORR:
                d["fwd_total"], d["rev_total"] = int(o.fwd_total), int(o.rev_total)
            res.append(d)
        return res

    # ---- the hot path
    def run_batch(self, jobs):
        """Host arrays in, per-shift int64 counters out (H2D + build + window + D2H)."""
        if not jobs:
            return []
        args = self._marshal(jobs)
        outs, arrays = self._outputs(jobs)
        self._check(self._L.pmb_run_batch(self._h, args, outs, len(jobs)))
        return self._collect(jobs, outs, arrays)

    # ---- the collective (include/pymasc_b200.h: pmb_comm_*, pmb_allreduce)
    COMM_ID_BYTES = 128

    @staticmethod
    def comm_get_id():
        """128 bytes that name a new communicator (rank 0 creates them, every rank gets a copy)."""
        buf = ctypes.create_string_buffer(Device.COMM_ID_BYTES)
        L = load_library()
        if L.pmb_comm_get_id(buf) != PMB_OK:
            raise DeviceError("pmb_comm_get_id failed: %s" % L.pmb_last_error(None).decode())
        return buf.raw

    def comm_init(self, comm_id, rank, world):
        buf = ctypes.create_string_buffer(bytes(comm_id), Device.COMM_ID_BYTES)
        self._check(self._L.pmb_comm_init(self._h, buf, int(rank), int(world)))
        self.comm_rank, self.comm_world = int(rank), int(world)

    @staticmethod
    def comm_init_all(devices):
        """One communicator over the contexts of one process (one per device)."""
        n = len(devices)
        arr = (ctypes.c_void_p * n)(*[d._h.value for d in devices])
        devices[0]._check(devices[0]._L.pmb_comm_init_all(arr, n))
        for i, d in enumerate(devices):
            d.comm_rank, d.comm_world = i, n

    @staticmethod
    def allreduce(batches):
        """In-place sum of the batches' device counter blocks over their contexts' communicator
        (one batch in a process per GPU; all of them in a single process)."""
        n = len(batches)
        ctxs = (ctypes.c_void_p * n)(*[b.dev._h.value for b in batches])
        bs = (ctypes.c_void_p * n)(*[b._h.value for b in batches])
        batches[0].dev._check(batches[0].dev._L.pmb_allreduce(ctxs, bs, n))

    NOALLOC_MAX_SHIFT = 4095

    @staticmethod
    def noalloc_out_int64(max_shift):
        return NCOUNTERS * (int(max_shift) + 1) + 3

    def run_chrom_noalloc(self, chrom_len, read_len, max_shift, flags, d_fwd, n_fwd, d_rev, n_rev,
                          d_begins, d_ends, n_intervals, d_out, out_int64, lo=0, hi=0):
        """pmb_run_chrom_noalloc: DEVICE pointers (integers) in, counters into the caller's device
        block; asynchronous on the context's stream, allocates nothing (see include/pymasc_b200.h)."""
        a = ChromArgs()
        a.chrom_len, a.read_len, a.max_shift, a.flags = int(chrom_len), int(read_len), int(max_shift), int(flags)
        a.fwd_pos, a.n_fwd = d_fwd or None, int(n_fwd)
        a.rev_pos, a.n_rev = d_rev or None, int(n_rev)
        a.map_begin, a.map_end, a.n_intervals = d_begins or None, d_ends or None, int(n_intervals)
        a.range_lo, a.range_hi = int(lo), int(hi)
        self._check(self._L.pmb_run_chrom_noalloc(self._h, ctypes.byref(a), ctypes.c_void_p(int(d_out)),
                                                  int(out_int64)))

    def prepare(self, jobs):
        return Batch(self, jobs)

    def timing(self):
        t = Timing()
        self._check(self._L.pmb_get_timing(self._h, ctypes.byref(t)))
        return {"build_ms": t.build_ms, "window_ms": t.window_ms, "total_ms": t.total_ms,
                "launches": t.launches, "window_launches": t.window_launches,
                "window_word_shift_pairs": t.window_word_shift_pairs, "path": t.path,
                "ev_fwd_ms": t.ev_fwd_ms, "ev_rev_ms": t.ev_rev_ms, "ev_edge_ms": t.ev_edge_ms}

    def pipe_probe(self, what):
        ops, ms = ctypes.c_double(), ctypes.c_float()
        self._check(self._L.pmb_pipe_probe(self._h, int(what), ctypes.byref(ops), ctypes.byref(ms)))
        return ops.value, ms.value


class Batch(object):
    """Inputs resident on the device: prepare once, execute (repeatedly), fetch."""

    def __init__(self, device, jobs):
        self.dev, self.jobs = device, list(jobs)
        args = device._marshal(self.jobs)
        h = ctypes.c_void_p()
        device._check(device._L.pmb_batch_prepare(device._h, args, len(self.jobs), ctypes.byref(h)))
        self._h = h

    def execute(self):
        self.dev._check(self.dev._L.pmb_batch_execute(self.dev._h, self._h))

    def allreduce(self):
        """Sum this batch's device counters over the ranks of the context's communicator."""
        Device.allreduce([self])

    def validate(self):
        """After the first execute, before the device counters are read without a fetch (an
        all-reduce): the list path's inputs turned out unsorted -> recompute on the event path."""
        self.dev._check(self.dev._L.pmb_batch_validate(self.dev._h, self._h))

    def fetch(self):
        outs, arrays = self.dev._outputs(self.jobs)
        self.dev._check(self.dev._L.pmb_batch_fetch(self.dev._h, self._h, outs))
        return self.dev._collect(self.jobs, outs, arrays)

    def counters(self):
        """(device pointer, number of int64, per-counter stride) of the counter block."""
        p, n, s = ctypes.c_void_p(), ctypes.c_uint64(), ctypes.c_uint32()
        self.dev._check(self.dev._L.pmb_batch_counters(self.dev._h, self._h, ctypes.byref(p),
                                                       ctypes.byref(n), ctypes.byref(s)))
        return p.value, n.value, s.value

    def free(self):
        if getattr(self, "_h", None) and getattr(self.dev, "_h", None):
            self.dev._L.pmb_batch_free(self.dev._h, self._h)
        self._h = None

    __del__ = free
