"""developer aid: small invocations of the kernels added last (FFJORD map, zero-padded generic shapes, fast weight-gradient
tiles, small-batch CTA sizes) for `compute-sanitizer --tool memcheck python tools/sanitize_new_paths.py`"""
import os, sys, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpflow_b200 as G
z = np.load(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "ffjord_dsprites_f32.npz"))
hid = [int(h) for h in z["hidden"]]
for dt in (torch.float32, torch.float64):
    m = G.FFJORDMap(int(z["d"]), hid, int(z["n_blocks"]), torch.from_numpy(z["params"]).to(dt), [float(t) for t in z["t_end"]], 2).cuda()
    for B in (1, 77, 300):
        x = torch.randn(B, 10, device="cuda", dtype=dt)
        y, ld = m(x); xb = m.backward(y)
        assert bool(torch.isfinite(y).all() and torch.isfinite(ld).all() and torch.isfinite(xb).all())
m = G.FFJORDMap(int(z["d"]), hid, int(z["n_blocks"]), torch.from_numpy(z["params"]), [float(t) for t in z["t_end"]], 1).cuda()
xb = m.backward(torch.randn(148 * 256 + 5, 10, device="cuda"))          # two samples per thread, ragged last tile
assert bool(torch.isfinite(xb).all())
for d, nb, hidden, B in [(10, 9, (50, 50, 50), 37), (3, 2, (64, 64), 150), (6, 3, (50, 100), 33), (20, 4, (128, 128), 40),
                         (5, 1, (7, 9), 33), (10, 9, (50, 50, 50), 700)]:
    cfg = G.FlowConfig(d=d, nb=nb, hidden=hidden, n_steps=2)
    params = 0.05 * torch.randn(cfg.n_params, device="cuda")
    x_map = torch.randn(B, d, device="cuda"); x_data = 1.1 * x_map.flip(-1) + 0.1
    res = G.fused_train_step(cfg, params, x_data, x_map, want_y=True)
    assert bool(torch.isfinite(res["grads"]).all() and torch.isfinite(res["y"]).all())
    y = G.flow_forward(cfg, params, x_map, wrap=True)
torch.cuda.synchronize()
print("sanitize_new_paths: ok")
