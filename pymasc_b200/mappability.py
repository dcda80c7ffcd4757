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
autocorrelation.  ``nworker`` only sizes the thread pool that decodes the BigWig (1, the
reference's default, means all cores).  There is no CPU fallback.
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

        # /path/to/track.bw -> /path/to/track_mappability.json unless told otherwise
        self.map_path = map_path or Path(path).with_name(Path(path).with_suffix("").name + "_mappability.json")
        self._adopt_cache_or_prepare_to_write()

    # ------------------------------------------------------------------ the JSON cache
    # Behaviour of handler/mappability.py:197-309: a readable, well-formed cache that covers at
    # least this max_shift and every chromosome of the BigWig is used (truncated to this
    # max_shift) and nothing is calculated or written; a cache for a shorter max_shift, a
    # malformed or an unreadable one is recalculated and overwritten; a cache path that is not a
    # file, or a target that cannot be written, is JSONIOError.
    def _adopt_cache_or_prepare_to_write(self):
        target = Path(self.map_path)
        if target.exists() and not target.is_file():
            logger.critical("Specified path is not file: '{}'".format(self.map_path))
            raise JSONIOError
        cached = self._usable_cache() if target.exists() else None
        if cached is not None:
            top = self.max_shift + 1
            self.mappable_len = cached["__whole__"][:top]
            self.chrom2mappable_len = {c: v[:top] for c, v in cached["references"].items()}
            self.chrom2is_called = dict.fromkeys(self.chromsizes, True)
            self.is_called, self.need_save_stats = True, False
            logger.info("Use mappability stats read from '{}'".format(self.map_path))
            return
        if target.exists():
            if not os.access(self.map_path, os.W_OK):
                logger.critical("Failed to overwrite '{}'".format(self.map_path))
                raise JSONIOError
            logger.warning("Existing file '{}' will be overwritten.".format(self.map_path))
        else:
            folder = str(target.parent) or "."
            try:
                os.makedirs(folder, exist_ok=True)
            except OSError:
                raise JSONIOError
            if not os.access(folder, os.W_OK):
                logger.critical("Directory '{}' is not writable.".format(folder))
                raise JSONIOError
        logger.info("Calcurate mappable length with max shift size {}.".format(self.max_shift))

    def _usable_cache(self):
        """The parsed cache file if it can stand in for a calculation, else None."""
        try:
            with open(self.map_path) as fh:
                stats = json.load(fh)
            cached_shift = stats["max_shift"]
            lists = [stats["__whole__"]] + [stats["references"][c] for c in self.chromsizes]
        except OSError as e:
            logger.error("Failed to read '{}': [Errno {}] {}".format(self.map_path, e.errno, e))
            return None
        except (TypeError, ValueError, KeyError, IndexError, OverflowError) as e:
            logger.error("Failed to load json file '{}': {!r}".format(self.map_path, e))
            return None
        if cached_shift < self.max_shift:
            logger.info("Specified shift length longer than former analysis. The stats will be updated.")
            return None
        if any(len(v) != cached_shift + 1 for v in lists):
            logger.error("Max shift length unmatched in '{}'.".format(self.map_path))
            return None
        return stats

    def save_mappability_stats(self):
        """Write the cache (keys max_shift / __whole__ / references, 4-space indent, sorted keys:
        the layout of handler/mappability.py:296-301) unless it was read from disk."""
        if not self.need_save_stats:
            return logger.info("Mappability stats updating is not required.")
        logger.info("Save mappable length to '{}'".format(self.map_path))
        doc = {"max_shift": int(self.max_shift), "__whole__": [int(x) for x in self.mappable_len],
               "references": {c: [int(x) for x in v] for c, v in self.chrom2mappable_len.items()}}
        try:
            with open(self.map_path, "w") as fh:
                json.dump(doc, fh, indent=4, sort_keys=True)
        except OSError as e:
            logger.error("Faild to output: {}\n[Errno {}] {}".format(e.filename, e.errno, e))
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
                if len(targets) > 1:
                    reader.prefetch(self.MAPPABILITY_THRESHOLD, self.nworker if self.nworker > 1 else 0)
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
