"""Latency of the per-step gradient all-reduce as the train step sees it: every rank runs a ~15 ms kernel, then
all_reduce(SUM) of n floats on the same stream; CUDA events give (a) the time from the end of the kernel to the end of the
all-reduce on each rank (max over ranks) and (b) the same with the kernel replaced by nothing (pure collective latency).
torchrun --nproc-per-node N tools/allreduce_latency.py [n_floats]"""
import os
import sys

import torch
import torch.distributed as dist

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = int(sys.argv[1]) if len(sys.argv) > 1 else 338
buf = torch.ones(n, device="cuda")
a = torch.randn(8192, 8192, device="cuda")
for with_kernel in (True, False):
    ts = []
    for it in range(25):
        e0, e1, e2 = (torch.cuda.Event(enable_timing=True) for _ in range(3))
        e0.record()
        if with_kernel:
            for _ in range(3):
                a @ a
        e1.record()
        dist.all_reduce(buf)
        e2.record()
        torch.cuda.synchronize()
        if it >= 5:
            ts.append((e0.elapsed_time(e1), e1.elapsed_time(e2)))
    t = torch.tensor([sum(x[0] for x in ts) / len(ts), sum(x[1] for x in ts) / len(ts), max(x[1] for x in ts)], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        print("world %d n=%d kernel=%s: kernel %.3f ms, all_reduce after it %.3f ms (mean, max over ranks), worst %.3f ms" % (
            world, n, with_kernel, float(t[0]), float(t[1]), float(t[2])))
dist.destroy_process_group()
