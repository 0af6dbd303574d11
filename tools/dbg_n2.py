"""developer aid: per-phase device timing of the data-parallel step under torchrun (not part of the product path)"""
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpflow_b200 as G
from gpflow_b200 import ops
from bench import synth_batch
rank = int(os.environ.get("RANK", 0)); lr = int(os.environ.get("LOCAL_RANK", 0)); world = int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
if world > 1: dist.init_process_group("nccl", device_id=dev)
B = (1 << 24) // world
cfg = G.FlowConfig(d=2, nb=1, hidden=(15, 15), n_steps=15)
torch.manual_seed(0); flat = (0.3 * torch.randn(cfg.n_params)).to(dev)
xd, xm = synth_batch(B, 2, 1234 + rank, dev)
buf = torch.empty(cfg.n_params + 3, device=dev)
mode = sys.argv[1] if len(sys.argv) > 1 else "both"
def step():
    ops.fused_train_step(cfg, flat, xd, xm, loss_scale=1.0 / (B * world), out=buf)
    if world > 1 and mode != "noar": dist.all_reduce(buf)
for _ in range(3): step()
torch.cuda.synchronize()
if world > 1: dist.barrier()
torch.cuda.synchronize()
evs = [torch.cuda.Event(enable_timing=True) for _ in range(11)]
evs[0].record()
for i in range(10):
    step(); evs[i + 1].record()
    if mode == "sync": torch.cuda.synchronize()
torch.cuda.synchronize()
print(rank, mode, ["%.1f" % evs[i].elapsed_time(evs[i + 1]) for i in range(10)], flush=True)
if world > 1: dist.destroy_process_group()
