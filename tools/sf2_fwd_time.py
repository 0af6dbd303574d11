"""Forward op (gp_flow: erf -> RK4 -> erfinv, log-det) of the C4 shape, 2^22 samples: ms per call, CUDA events."""
import os, sys, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "oracle"))
import gpflow_b200 as G
import gpflow_oracle as O
spec = O.FlowSpec(d=2, nb=1, hidden=[15, 15], n_steps=15)
cfg = G.FlowConfig(d=2, nb=1, hidden=(15, 15), n_steps=15)
flat = O.flatten_params(O.init_params(spec, 0.3, 1, torch.float32)).cuda()
x = torch.randn(1 << 22, 2, device="cuda").clamp(-3, 3)
for lag in (False, True):
    for _ in range(3):
        G.flow_forward(cfg, flat, x, wrap=True, lagrangian=lag)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        out = G.flow_forward(cfg, flat, x, wrap=True, lagrangian=lag)
    e1.record(); torch.cuda.synchronize()
    print(G.kernel_variant(cfg, op=0, lagrangian=lag), "lagrangian=%s  %.3f ms per 2^22 samples  %.3g samples/s" % (lag, e0.elapsed_time(e1) / 5, (1 << 22) / (e0.elapsed_time(e1) / 5e3)))
