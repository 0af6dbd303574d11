"""C3 shape WITH the Euler penalization on the generic kernel: GPFLOW_GENERIC_ALIGN=0/1 (set in the environment), B samples."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))
import gpflow_oracle as O
import gpflow_b200 as G
B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
spec = O.FlowSpec(d=10, nb=9, hidden=[50, 50, 50], n_steps=15)
flat = O.flatten_params(O.init_params(spec, 0.05, 0, torch.float32)).cuda()
cfg = G.FlowConfig(d=10, nb=9, hidden=(50, 50, 50), n_steps=15)
g = torch.Generator(device="cuda").manual_seed(1)
xm = torch.randn(B, 10, generator=g, device="cuda"); xd = 0.9 * xm.roll(1, 1) + 0.1
zw = torch.randn(15, 2, 10, generator=g, device="cuda")
for i in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); r = G.fused_train_step(cfg, flat, xd, xm, euler_coef=5e-4, zw=zw); e1.record(); e1.synchronize()
print("align=%s B=%d: %.2f ms -> %.4g samples/s  ot %.6f reg %.6e |g| %.6e" % (os.environ.get("GPFLOW_GENERIC_ALIGN"), B, e0.elapsed_time(e1),
      B / e0.elapsed_time(e1) * 1e3, float(r["scalars"][0]), float(r["scalars"][1]), float(r["grads"].norm())))
