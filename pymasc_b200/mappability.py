"""Mappable-length statistics of a BigWig track on the B200.

``MappabilityHandler`` mirrors PyMaSC/handler/mappability.py:99-361 (which extends
``MappableLengthCalculator``, PyMaSC/core/mappability.pyx:39-215): per chromosome the number of
position pairs ``sum_j M[j] M[j+s]`` for ``s = 0 .. max_shift`` over the track's mappable
positions, their genome-wide sum, and the ``*_mappability.json`` cache (keys ``max_shift``,
``__whole__``, ``references``) that ``pymasc`` and ``pymasc-precalc`` read and write.

The reference streams the intervals through a sliding window of ``max_shift`` positions per
chromosome (``_feed_track``, mappability.pyx:176-205) on ``nworker`` processes.  Here every
chromosome that still has to be calculated becomes one ``PMB_FLAG_AUTOCORR`` job
(include/pymasc_b200.h) and all of them run as ONE device batch: the track is gathered into bit
arrays and the same run-edge kernels that produce ``popc(D_k)`` for MSCC count the
autocorrelation.  ``nworker`` is accepted and ignored.  There is no CPU fallback.
"""
import json
import logging
import os
from pathlib import Path

import numpy as np

from . import _shim
from .io_native import BigWigReader

logger = logging.getLogger(__name__)


class BWIOError(IOError):
    pass


class JSONIOError(IOError):
    pass


class NeedUpdate(Exception):
    pass


class MappabilityHandler(object):
    MAPPABILITY_THRESHOLD = 1.

    @staticmethod
    def calc_mappable_len_required_shift_size(readlen, max_shift):
        # handler/mappability.py:120-135
        return max_shift - readlen + 1 if max_shift > 2 * readlen - 1 else readlen

    def __init__(self, path, max_shift=0, readlen=0, map_path=None, nworker=1, engine=None):
        max_shift = self.calc_mappable_len_required_shift_size(readlen, max_shift)
        self.nworker = nworker
        path_str = os.fspath(path)
        if not os.access(path_str, os.R_OK):
            reason = "file is unreadable." if Path(path).is_file() else "no such file."
            logger.critical("Failed to open '{}': {}".format(path, reason))
            raise BWIOError
        self.path = path_str
        reader = BigWigReader(path_str)
        self.chromsizes = dict(reader.chromsizes)
        reader.close()
        self.max_shift = max_shift
        self.readlen = readlen
        self.chrom2is_called = {c: False for c in self.chromsizes}
        self.chrom2mappable_len = {}
        self.mappable_len = [0] * (max_shift + 1)
        self.is_called = False
        self.need_save_stats = True
        self._engine = engine

        if map_path:
            self.map_path = map_path
        else:       # /path/to/file.bw -> /path/to/file_mappability.json
            p = Path(path)
            self.map_path = p.parent / (p.with_suffix("").name + "_mappability.json")

        if not Path(self.map_path).exists():
            self._check_saving_directory_is_writable()
            logger.info("Calcurate mappable length with max shift size {}.".format(max_shift))
        elif not Path(self.map_path).is_file():
            logger.critical("Specified path is not file: '{}'".format(self.map_path))
            raise JSONIOError
        elif not os.access(self.map_path, os.R_OK):
            logger.error("Failed to read '{}'".format(self.map_path))
        else:
            self._try_load_mappability_stats()
            if self.need_save_stats:
                self._check_stats_is_overwritable()
                logger.info("Calcurate mappable length with max shift size {}.".format(max_shift))
            else:
                logger.info("Use mappability stats read from '{}'".format(self.map_path))

    # ------------------------------------------------------------------ the JSON cache
    def _check_saving_directory_is_writable(self):
        dirname = str(Path(self.map_path).parent) or '.'
        try:
            os.makedirs(dirname, exist_ok=True)
        except OSError:
            raise JSONIOError
        if not os.access(dirname, os.W_OK):
            logger.critical("Directory '{}' is not writable.".format(dirname))
            raise JSONIOError

    def _try_load_mappability_stats(self):
        try:
            stats = self._read_mappability_stats()
        except IOError as e:
            logger.error("Failed to read '{}'".format(self.map_path))
            logger.error("[Errno {}] {}".format(e.errno, str(e)))
        except (TypeError, OverflowError, ValueError, KeyError, IndexError):
            logger.error("Failed to load json file: '{}'".format(self.map_path))
        except NeedUpdate:
            pass
        else:
            self._load_mappability_stats(stats)

    def _read_mappability_stats(self):
        with open(self.map_path) as f:
            stats = json.load(f)
        for k in ("max_shift", "__whole__", "references"):
            if k not in stats:
                logger.error("Mandatory key '{}' not found.".format(k))
                raise KeyError(k)
        if stats["max_shift"] < self.max_shift:
            logger.info("Specified shift length longer than former analysis. The stats will be updated.")
            raise NeedUpdate
        if stats["max_shift"] != len(stats["__whole__"]) - 1:
            logger.error("Max shift length for whole genome unmatched.")
            raise IndexError
        for ref in self.chromsizes:
            if ref not in stats["references"]:
                logger.error("Reference '{}' not found.".format(ref))
                raise KeyError(ref)
            if stats["max_shift"] != len(stats["references"][ref]) - 1:
                logger.error("Max shift length for '{}' unmatched.".format(ref))
                raise IndexError
        return stats

    def _load_mappability_stats(self, stats):
        self.mappable_len = stats["__whole__"][:self.max_shift + 1]
        self.chrom2mappable_len = {ref: b[:self.max_shift + 1] for ref, b in stats["references"].items()}
        self.chrom2is_called = {ref: True for ref in self.chromsizes}
        self.is_called = True
        self.need_save_stats = False

    def _check_stats_is_overwritable(self):
        if not os.access(self.map_path, os.W_OK):
            logger.critical("Failed to overwrite '{}'".format(self.map_path))
            raise JSONIOError
        logger.warning("Existing file '{}' will be overwritten.".format(self.map_path))

    def save_mappability_stats(self):
        # handler/mappability.py:283-309: same keys, indent and key order
        if not self.need_save_stats:
            return logger.info("Mappability stats updating is not required.")
        logger.info("Save mappable length to '{}'".format(self.map_path))
        try:
            with open(self.map_path, 'w') as f:
                json.dump({
                    "max_shift": int(self.max_shift),
                    "__whole__": [int(x) for x in self.mappable_len],
                    "references": {c: [int(x) for x in v] for c, v in self.chrom2mappable_len.items()},
                }, f, indent=4, sort_keys=True)
        except IOError as e:
            logger.error("Faild to output: {}\n[Errno {}] {}".format(e.filename, e.errno, str(e)))
        self.need_save_stats = False

    # ------------------------------------------------------------------ the calculation
    def _run(self, jobs):
        if self._engine is None:
            from .calculator import DeviceEngine
            self._engine = DeviceEngine()
        return self._engine.run(jobs)

    def calc_mappability(self, chrom=None):
        """Every chromosome not calculated yet (or just `chrom`), one device batch."""
        if chrom is None:
            targets = [c for c, done in self.chrom2is_called.items() if not done]
        elif self.chrom2is_called[chrom]:
            return None
        else:
            targets = [chrom]
        if targets:
            reader = BigWigReader(self.path)
            rl = max(int(self.readlen), 2)      # only sizes the bit array; the counts do not depend on it
            jobs = []
            try:
                for c in targets:
                    logger.info("Calc {} mappable length...".format(c))
                    b, e = reader.fetch_arrays(self.MAPPABILITY_THRESHOLD, c)
                    jobs.append(_shim.Job(max(self.chromsizes[c], int(e.max()) if e.size else 0), rl,
                                          self.max_shift, _shim.FLAG_AUTOCORR, None, None, b, e))
            finally:
                reader.close()
            for c, out in zip(targets, self._run(jobs)):
                self.chrom2mappable_len[c] = tuple(int(x) for x in out["dmp"])
                self.chrom2is_called[c] = True
        if all(self.chrom2is_called.values()):
            self.is_called = True
        if chrom is None or self.is_called:
            self._sumup_mappability()

    def _sumup_mappability(self):
        total = np.zeros(self.max_shift + 1, dtype=np.int64)
        for length in self.chrom2mappable_len.values():
            total += np.asarray(length[:self.max_shift + 1], dtype=np.int64)
        self.mappable_len = [int(x) for x in total]

    def get_mappable_len(self, chrom=None, shift_from=None, shift_to=None, force=False):
        # core/mappability.pyx:207-241
        if chrom is not None:
            if chrom not in self.chrom2is_called:
                return None
            if self.chrom2is_called[chrom]:
                return self.chrom2mappable_len[chrom][shift_from:shift_to]
            elif force:
                self.calc_mappability(chrom)
                return self.chrom2mappable_len[chrom][shift_from:shift_to]
            raise KeyError("Mappable length for '{}' is not calculated yet.".format(chrom))
        if not self.is_called:
            self.calc_mappability()
        return self.mappable_len[shift_from:shift_to]

    def close(self):
        pass
