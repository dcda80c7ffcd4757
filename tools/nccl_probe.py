"""All-reduce latency of small int64 blocks on this box (torchrun, one rank per GPU)."""
import os, time, torch, torch.distributed as dist
local = int(os.environ["LOCAL_RANK"]); torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
for n, dt in ((123000, torch.int64), (123000, torch.float32), (12300, torch.int64), (1 << 20, torch.int64)):
    t = torch.ones(n, dtype=dt, device="cuda")
    for _ in range(20): dist.all_reduce(t)
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); e0.record()
    for _ in range(200): dist.all_reduce(t)
    e1.record(); torch.cuda.synchronize(); w = time.perf_counter() - t0
    if dist.get_rank() == 0:
        print("n=%d %s: %.1f us/call device, %.1f us/call wall" % (n, dt, e0.elapsed_time(e1) * 5, w * 5e3), flush=True)
dist.destroy_process_group()
