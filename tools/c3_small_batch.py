"""developer aid: device time of one fused train step of the generic kernel at the reference's real batch (B = 1024, C3)"""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpflow_b200 as G
for d, nb, hidden in [(10, 9, (50, 50, 50)), (64, 63, (256, 256))][: int(os.environ.get("SB_SHAPES", "2"))]:
    cfg = G.FlowConfig(d=d, nb=nb, hidden=hidden, n_steps=15)
    params = 0.03 * torch.randn(cfg.n_params, device="cuda")
    for B in (256, 1024, 2048):
        x_map = torch.randn(B, d, device="cuda"); x_data = 1.1 * x_map.flip(-1) + 0.1
        for _ in range(2): G.fused_train_step(cfg, params, x_data, x_map)
        ts = []
        for _ in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); G.fused_train_step(cfg, params, x_data, x_map); e1.record(); e1.synchronize(); ts.append(e0.elapsed_time(e1))
        ts.sort()
        print("WPC=%s d=%d nb=%d H=%s B=%d: %.2f ms/step -> %.3g samples/s" % (os.environ.get("GPFLOW_GENERIC_WPC", "auto"), d, nb, hidden, B, ts[1], B / ts[1] * 1e3), flush=True)
