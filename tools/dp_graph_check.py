"""Two ranks under torchrun: the graph-replayed tensor-core train step inside a data-parallel loop (NCCL process group alive
during the capture) against eager launches -- same losses and parameters bit for bit on every rank."""
import os, sys, torch
import torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import gpflow_b200 as G
rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl")
dev = torch.device("cuda", local)
out = {}
for mode in ("1", "0"):
    os.environ["GPFLOW_TC_GRAPH"] = mode
    torch.manual_seed(0)
    flow = G.get_gp_flow(10, {"nb_neurons": [50, 50, 50], "nb_div_blocks": 9}, 3).to(dev)
    with torch.no_grad():
        for p in flow.parameters():
            p.normal_(0.0, 0.03)
    net = G.GPFlowNetwork(flow, map_module=None)
    opt = net.configure_fused_optimizer(lr=2e-3, weight_decay=1e-4)
    g = torch.Generator(device=dev).manual_seed(3 + rank)
    losses = []
    for i in range(4):
        x_map = torch.randn(3072, 10, generator=g, device=dev)
        loss, _ = net.fused_training_step(0.9 * x_map.roll(1, 1) + 0.1, x_map)
        opt.step()
        losses.append(float(loss))
    torch.cuda.synchronize()
    out[mode] = (losses, opt.flat.clone())
same = out["1"][0] == out["0"][0] and torch.equal(out["1"][1], out["0"][1])
print("rank %d: graph == eager: %s  losses %s" % (rank, same, ["%.6f" % v for v in out["1"][0]]), flush=True)
dist.barrier()
dist.destroy_process_group()
assert same
