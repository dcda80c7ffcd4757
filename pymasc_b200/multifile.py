"""Several read sets over ONE mappability track in one device batch (BASELINE config 5).

A multi-file ``pymasc`` run (PyMaSC/pymasc.py loops over its input files, one CalcHandler
each) loads the same BigWig for every file and rebuilds every chromosome's mappability bit
array, and the per-shift mappable lengths popc(D_k) (mscc.pyx:291-298), once per file although
neither depends on the reads.  Here the calculators of all files are finished together: jobs of
one chromosome that read the same track form a run in which only the first job carries the
intervals and the others are flagged ``FLAG_SAME_TRACK`` (include/pymasc_b200.h), so M, its
derived arrays and dmp are built once per chromosome and the read-less chromosomes'
autocorrelation (mscc.pyx:207-215) runs once instead of once per file.  Results are identical
to finishing every calculator on its own.
"""
from . import _shim


def _shared_copy(job):
    """`job` as a FLAG_SAME_TRACK follower: its reads, no intervals of its own."""
    j = _shim.Job(job.chrom_len, job.read_len, job.max_shift, job.flags | _shim.FLAG_SAME_TRACK,
                  job.fwd_pos, job.rev_pos, None, None, job.lo, job.hi)
    j.track_key = job.track_key
    return j


def merge_shared_tracks(job_lists):
    """Merge the job lists of several calculators into one batch.

    Jobs with equal ``track_key`` and geometry become one contiguous run, first job unchanged
    (it owns the track), the rest as FLAG_SAME_TRACK followers; equal autocorrelation jobs
    collapse into one.  Jobs without a track keep their place.  Returns (merged, route) with
    ``route[f][i]`` = index in `merged` whose output answers job i of list f."""
    groups, order = {}, []
    for f, jobs in enumerate(job_lists):
        for i, j in enumerate(jobs):
            shareable = j.track_key is not None and j.flags & (_shim.FLAG_MSCC | _shim.FLAG_AUTOCORR)
            key = ((j.track_key, j.chrom_len, j.read_len, j.max_shift, j.lo, j.hi,
                    bool(j.flags & _shim.FLAG_AUTOCORR)) if shareable else ("solo", f, i))
            if key not in groups:
                groups[key] = []
                order.append(key)
            groups[key].append((f, i))
    merged = []
    route = [[None] * len(jobs) for jobs in job_lists]
    for key in order:
        members = groups[key]
        f0, i0 = members[0]
        first = job_lists[f0][i0]
        if first.flags & _shim.FLAG_AUTOCORR:
            for f, i in members:
                route[f][i] = len(merged)
            merged.append(first)
            continue
        for n, (f, i) in enumerate(members):
            route[f][i] = len(merged)
            merged.append(job_lists[f][i] if n == 0 else _shared_copy(job_lists[f][i]))
    return merged, route


class MultiFileCalculation(object):
    """Finishes several ``CCBitArrayCalculator`` objects -- one per input file, all constructed
    with the SAME ``bwfeeder`` object -- in one device batch per (read_len, max_shift).

    Feed each calculator as usual, then call ``finishup_calculation()`` here instead of on the
    calculators; afterwards ``get_result`` / ``get_whole_result`` of each calculator answer as
    if it had been finished alone."""

    def __init__(self, calculators, engine=None):
        self.calculators = list(calculators)
        self._engine = engine

    def finishup_calculation(self):
        by_geometry = {}
        for c in self.calculators:
            by_geometry.setdefault((c.read_len, c.max_shift), []).append(c)
        for calcs in by_geometry.values():
            collected = []
            for c in calcs:
                if c._chr != '' and not c._buff_flashed:
                    c._close_chromosome()
                collected.append(c._collect_jobs(fill=[c._chr] + list(c.references)))
                c._buff_flashed = True
            merged, route = merge_shared_tracks([jobs for jobs, _ in collected])
            engine = self._engine if self._engine is not None else calcs[0]._engine
            outs = engine.run(merged) if merged else []
            for c, (jobs, fills), r in zip(calcs, collected, route):
                c._apply_results([outs[k] for k in r], fills)
