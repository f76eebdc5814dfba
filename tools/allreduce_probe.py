#!/usr/bin/env python
"""How long does the histogram all-reduce take by itself?  torchrun --nproc-per-node N tools/allreduce_probe.py"""
import os
import torch
import torch.distributed as dist

local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
for nbytes, dt in ((4 << 20, torch.int32), (8 << 20, torch.int64), (8 << 20, torch.int32), (64 << 20, torch.int64)):
    x = torch.ones(nbytes // dt.itemsize, dtype=dt, device="cuda")
    for _ in range(10):
        dist.all_reduce(x)
    torch.cuda.synchronize()
    dist.barrier()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(50)]
    for a, b in ev:
        a.record()
        dist.all_reduce(x)
        b.record()
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) for a, b in ev)
    # with a host sync between the calls, as in the bench step
    hs = []
    for _ in range(30):
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); dist.all_reduce(x); b.record(); torch.cuda.synchronize()
        hs.append(a.elapsed_time(b))
    hs.sort()
    if dist.get_rank() == 0:
        print(f"all_reduce {nbytes >> 20} MiB {dt}: back-to-back median {ts[len(ts)//2]*1e3:.1f} us, min {ts[0]*1e3:.1f} us; "
              f"with host sync median {hs[len(hs)//2]*1e3:.1f} us", flush=True)
dist.destroy_process_group()
