"""developer aid: device time of one fused train step at the reference's real batch size (B = 1024, C1/C2)"""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpflow_b200 as G
from gpflow_b200 import ops
cfg = G.FlowConfig(d=2, nb=1, hidden=(15, 15), n_steps=15)
torch.manual_seed(0)
flat = (0.3 * torch.randn(cfg.n_params)).cuda()
for B in (1024, 4096, 65536):
    xm = torch.randn(B, 2, device="cuda"); xd = 1.2 * xm.flip(-1) + 0.3
    for name, kw in (("sf2", {}), ("v1", {"force_v1": True}), ("v1+euler", {"force_v1": True, "euler_coef": 5e-4, "zw": torch.randn(15, 2, 2)}),
                     ("sf2e euler", {"euler_coef": 5e-4, "zw": torch.randn(15, 2, 2)})):
        for _ in range(5): ops.fused_train_step(cfg, flat, xd, xm, **kw)
        ts = []
        for _ in range(20):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); ops.fused_train_step(cfg, flat, xd, xm, **kw); e1.record(); e1.synchronize()
            ts.append(e0.elapsed_time(e1))
        ts.sort()
        print("B=%6d %-10s median %.3f ms  -> %.3g samples/s" % (B, name, ts[10], B / ts[10] * 1e3), flush=True)

# whole optimisation step at B = 1024 (C1): module API, wall clock over 300 steps (host launch overhead included)
import time
for name in ("torch.optim.Adam + .grad copies", "FlatAdam (flat buffers, one launch)"):
    torch.manual_seed(0)
    flow = G.get_gp_flow(2, {"nb_neurons": [15, 15]}, 15).cuda()
    with torch.no_grad():
        for p in flow.parameters():
            p.normal_(0.0, 0.3)
    net = G.GPFlowNetwork(flow)
    opt = torch.optim.Adam(flow.parameters(), lr=2e-3, weight_decay=1e-4) if name.startswith("torch") else \
        net.configure_fused_optimizer(lr=2e-3, weight_decay=1e-4)
    xm = torch.randn(1024, 2, device="cuda"); xd = 1.2 * xm.flip(-1) + 0.3
    for _ in range(20):
        net.fused_training_step(xd, xm); opt.step()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(300):
        net.fused_training_step(xd, xm); opt.step()
    torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 300
    print("B=1024 full step, %-38s %.3f ms/step -> %.3g samples/s" % (name, dt * 1e3, 1024 / dt), flush=True)
