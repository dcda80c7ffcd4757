"""Where does the all-reduce time of bench.py --gpus N go?  (torchrun, small workload)"""
import os, sys, time, torch, torch.distributed as dist, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from pymasc_b200 import _shim, sharding, synth
from pymasc_b200.distributed import counters_tensor
local = int(os.environ["LOCAL_RANK"]); torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
rank, world = dist.get_rank(), dist.get_world_size()
wl = synth.workload(n_reads=2_000_000, chroms=["chr%d" % i for i in range(1, 23)] + ["chrX", "chrY"], scale=0.1)
lengths = [w["length"] for w in wl]
jobs = sharding.jobs_for_rank(wl, sharding.plan(lengths, 36, 1000, world)[rank], 36, 1000)
dev = _shim.Device(local)
stream = torch.cuda.Stream()
dev.set_stream(stream.cuda_stream)
def timed(fn, n=50):
    for _ in range(5): fn()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(n): fn()
    e1.record(stream); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
with torch.cuda.stream(stream):
    b = dev.prepare(jobs)
    ctr = counters_tensor(b)
    own = torch.zeros_like(ctr)
    res = {}
    res["execute"] = timed(lambda: b.execute())
    res["allreduce(ctr)"] = timed(lambda: dist.all_reduce(ctr))
    res["allreduce(torch tensor)"] = timed(lambda: dist.all_reduce(own))
    res["execute+allreduce(ctr)"] = timed(lambda: (b.execute(), dist.all_reduce(ctr)))
    res["execute+copy+allreduce(own)"] = timed(lambda: (b.execute(), own.copy_(ctr), dist.all_reduce(own)))
    if rank == 0:
        print("numel", ctr.numel(), {k: round(v, 4) for k, v in res.items()}, flush=True)
    b.free()
dev.set_stream(None); dev.close()
dist.destroy_process_group()
