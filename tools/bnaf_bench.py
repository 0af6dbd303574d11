"""developer aid: device time of the frozen BNAF map forward (gpflow_bnaf_forward) at C4-sized batches"""
import os, sys, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpflow_b200 as G
z = np.load(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "bnaf_eight_gaussians_f32.npz"))
m = G.BNAFMap(int(z["dim"]), int(z["n_neurons"]), int(z["n_layers"]), torch.from_numpy(z["params"])).cuda()
for lb in (10, 16, 20, 24):
    B = 1 << lb
    x = 2.0 * torch.randn(B, 2, device="cuda")
    for _ in range(3): m(x)
    ts = []
    for _ in range(10):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); m(x); e1.record(); e1.synchronize(); ts.append(e0.elapsed_time(e1))
    ts.sort()
    print("B=2^%d  median %.3f ms -> %.3g samples/s" % (lb, ts[5], B / ts[5] * 1e3), flush=True)
