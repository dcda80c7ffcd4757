"""Where the end-to-end step goes: host wall time of prepare / execute / fetch / free on the
bench workload, for copy-then-compute and for the pipelined input copy, next to the bare
PCIe time of the same bytes.  Run on the GPU box: python tools/e2e_breakdown.py [world]
(world > 1: rank 0's share of the `world` position ranges -- the per-rank fixed cost of an N-GPU step)"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    import torch
    from pymasc_b200 import _shim, sharding, synth
    lengths = [L for _, L in synth.HG38]
    world = int(sys.argv[1]) if len(sys.argv) > 1 else 1
    pieces = sharding.plan(lengths, bench.READ_LEN, bench.MAX_SHIFT, world)[0]
    workload, _ = bench.build_workload("both")
    jobs = sharding.jobs_for_rank(workload, pieces, bench.READ_LEN, bench.MAX_SHIFT)
    nbytes = bench.pinned_pack(jobs)
    dev = _shim.Device(0)
    out = {"h2d_bytes": nbytes, "world": world}

    # bare PCIe: one pinned buffer of the same size
    src = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    dst = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    for _ in range(3):
        dst.copy_(src, non_blocking=True)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(10):
        dst.copy_(src, non_blocking=True)
    torch.cuda.synchronize()
    out["bare_h2d_ms"] = (time.perf_counter() - t0) * 100

    steps = 10
    for groups in (0, 4):
        dev.set_option("async_h2d", 1 if groups else 0)
        if groups:
            dev.set_option("h2d_groups", groups)
        acc = np.zeros(5)
        for it in range(steps + 3):
            t = [time.perf_counter()]
            b = dev.prepare(jobs); t.append(time.perf_counter())
            b.execute(); t.append(time.perf_counter())
            dev.synchronize(); t.append(time.perf_counter())
            b.fetch(); t.append(time.perf_counter())
            b.free(); t.append(time.perf_counter())
            if it >= 3:
                acc += np.diff(t)
        out["groups_%d" % groups] = dict(zip(
            ("prepare_ms", "execute_call_ms", "wait_ms", "fetch_ms", "free_ms"),
            (acc / steps * 1e3).round(3).tolist()))
        t0 = time.perf_counter()
        for _ in range(steps):
            dev.run_batch(jobs)
        out["groups_%d" % groups]["run_batch_ms"] = round((time.perf_counter() - t0) / steps * 1e3, 3)
    dev.set_option("async_h2d", 0)
    t0 = time.perf_counter()
    for _ in range(100):
        dev._marshal(jobs)
    out["marshal_ms"] = round((time.perf_counter() - t0) * 10, 3)
    out["device_timing_last"] = dev.timing()
    print(json.dumps(out))


if __name__ == "__main__":
    main()
