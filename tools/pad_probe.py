"""developer aid: train-step time of the generic kernel at the C3 shape against zero-padded hidden widths"""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpflow_b200 as G
B = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
for d, nb, hidden in [(10, 9, (50, 50, 50)), (10, 9, (52, 52, 52)), (10, 9, (56, 56, 56)), (10, 9, (64, 64, 64)), (10, 9, (128, 128, 128)), (12, 9, (64, 64, 64))]:
    cfg = G.FlowConfig(d=d, nb=nb, hidden=hidden, n_steps=15)
    n = [d + 1] + list(hidden) + [d * nb]
    P = sum(n[i] * n[i + 1] + n[i + 1] for i in range(len(n) - 1))
    params = 0.05 * torch.randn(P, device="cuda")
    x_map = torch.randn(B, d, device="cuda"); x_data = 1.1 * x_map.flip(-1) + 0.1
    for _ in range(2): G.fused_train_step(cfg, params, x_data, x_map)
    ts = []
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); G.fused_train_step(cfg, params, x_data, x_map); e1.record(); e1.synchronize(); ts.append(e0.elapsed_time(e1))
    ts.sort()
    print("d=%d nb=%d H=%s B=%d: %.1f ms -> %.3g samples/s (%s)" % (d, nb, hidden, B, ts[1], B / ts[1] * 1e3, G.kernel_variant(cfg)), flush=True)
