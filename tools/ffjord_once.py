"""developer aid: a few launches of gpflow_ffjord_integrate (backward, value only) for ncu"""
import os, sys, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpflow_b200 as G
z = np.load(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "ffjord_dsprites_f32.npz"))
m = G.FFJORDMap(int(z["d"]), [int(h) for h in z["hidden"]], int(z["n_blocks"]), torch.from_numpy(z["params"]),
                [float(t) for t in z["t_end"]], int(sys.argv[2]) if len(sys.argv) > 2 else 16).cuda()
x = torch.randn(int(sys.argv[1]) if len(sys.argv) > 1 else 37888, int(z["d"]), device="cuda")
for _ in range(3):
    y = m.backward(x)
torch.cuda.synchronize()
print("ok", float(y.abs().max()))
