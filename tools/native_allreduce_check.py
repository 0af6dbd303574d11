"""Two (or more) ranks under torchrun: the C ABI's all-reduce (gpflow_allreduce_grads, NCCL bound at run time) against
torch.distributed.all_reduce on the fused [grads | scalars] buffer of a train step -- same sums bit for bit (both are NCCL
ring / tree sums of the same inputs; compared at 1e-6), and the time of each per call.
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/native_allreduce_check.py"""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import gpflow_oracle as O  # noqa: E402
import gpflow_b200 as G  # noqa: E402
from gpflow_b200.data_parallel import NativeAllReduce  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl")
dev = torch.device("cuda", local)
for name, d, nb, hidden, B in (("c4 shape", 2, 1, (15, 15), 1 << 18), ("c5 shape", 64, 63, (256, 256), 1 << 11)):
    spec = O.FlowSpec(d=d, nb=nb, hidden=list(hidden), n_steps=15)
    flat = O.flatten_params(O.init_params(spec, 0.3 if d == 2 else 0.02, 0, torch.float32)).to(dev)
    g = torch.Generator(device=dev).manual_seed(100 + rank)
    x_map = torch.randn(B, d, generator=g, device=dev)
    x_data = 0.9 * x_map.roll(1, 1) + 0.1
    cfg = G.FlowConfig(d=d, nb=nb, hidden=hidden, n_steps=15)
    buf = G.fused_train_step(cfg, flat, x_data, x_map, loss_scale=1.0 / (B * world))["buf"].clone()
    a, b = buf.clone(), buf.clone()
    ar = NativeAllReduce(dev)
    ar.allreduce_(a)
    dist.all_reduce(b)
    torch.cuda.synchronize()
    rel = float((a - b).norm() / b.norm())
    times = {}
    for label, fn in (("gpflow_allreduce_grads", lambda t: ar.allreduce_(t)), ("torch.distributed.all_reduce", lambda t: dist.all_reduce(t))):
        t = buf.clone()
        for _ in range(5):
            fn(t)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        dist.barrier()
        e0.record()
        for _ in range(50):
            fn(t)
        e1.record()
        e1.synchronize()
        times[label] = e0.elapsed_time(e1) / 50 * 1e3
    if rank == 0:
        print("%s: %d floats, world %d: rel diff %.2e | %s" % (name, buf.numel(), world, rel,
              ", ".join("%s %.1f us" % kv for kv in times.items())), flush=True)
    assert rel < 1e-6
    ar.close()
dist.destroy_process_group()
