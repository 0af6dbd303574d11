"""developer aid: small invocations of the kernels added in round 2 (sf2r train / backward / forward-only, sf2e, potentials,
the no-allocation velocity, the wide finalize) for
    compute-sanitizer --tool memcheck  python tools/sanitize_round2.py
    compute-sanitizer --tool racecheck python tools/sanitize_round2.py      (shared-memory hazards between the lanes of a warp)
Batches are chosen to give ragged tiles, warps without a tile, and warps with two tiles (stash buffers wrap)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpflow_b200 as G
from gpflow_b200 import ops_aux
n_steps = int(os.environ.get("SAN_STEPS", "3"))
cfg = G.FlowConfig(d=2, nb=1, hidden=(15, 15), n_steps=n_steps)
g = torch.Generator(device="cuda").manual_seed(0)
params = 0.3 * torch.randn(cfg.n_params, device="cuda", generator=g)
for B in [int(b) for b in os.environ.get("SAN_BATCHES", "1,33,700").split(",")]:
    x_map = torch.randn(B, 2, device="cuda", generator=g).clamp(-3, 3)
    x_data = 1.2 * x_map.flip(-1) + 0.3
    zw = torch.randn(n_steps, 2, 2, device="cuda", generator=g)
    r = G.fused_train_step(cfg, params, x_data, x_map, want_y=True, want_grad_x=True)                       # sf2r
    assert bool(torch.isfinite(r["buf"]).all() and torch.isfinite(r["grad_x"]).all())
    r = G.fused_train_step(cfg, params, x_data, x_map, euler_coef=0.3, zw=zw, want_y=True, want_grad_x=True)   # sf2e
    assert bool(torch.isfinite(r["buf"]).all() and torch.isfinite(r["grad_x"]).all())
    gr, gx = G.flow_backward(cfg, params, x_map, torch.ones_like(x_map), wrap=True, training=True, euler_coef=0.3, zw=zw,
                             reg_cotangent=torch.tensor(0.5, device="cuda"))
    assert bool(torch.isfinite(gr).all() and torch.isfinite(gx).all())
    out = G.flow_forward(cfg, params, x_map, wrap=True, lagrangian=True)                                     # sf2r forward-only
    out = G.flow_forward(cfg, params, x_map, wrap=True, reverse=True)
vf = G.VelocityField(4, [5, 5, 5], nb_div_blocks=2).cuda()
f, J = ops_aux.potentials(vf.v, torch.rand(9, 5, device="cuda"))
assert bool(torch.isfinite(f).all() and torch.isfinite(J).all())
c1 = G.FlowConfig(d=5, nb=3, hidden=(5, 5, 5), n_steps=1)
x = torch.rand(40, 5, device="cuda") - 0.5
out = torch.empty_like(x)
ops_aux.velocity_static(c1, 0.3 * torch.randn(c1.n_params, device="cuda"), x, 0.3, out)
assert bool(torch.isfinite(out).all())
torch.cuda.synchronize()
print("sanitize_round2: ok")
