"""2-D nets NARROWER than the presets' [15,15] (d = 2, nb = 1, two hidden layers of width <= 15, fp32) run on the [15,15]
stream-function kernels with zero-padded parameters (gpflow_api.cu: pad2d_match / pad2d_run): a padded unit has h = tanh 0 = 0
and only zero weights leave it, so outputs and kept gradient entries are those of the narrow net.  Checked against the fp64
oracle (vector_field.py:44-128, velocity_field.py:20-79 restated) at the 1e-5 bar and against the generic kernel."""
import pytest
import torch

import gpflow_oracle as O
from golden_util import rel

pytestmark = pytest.mark.gpu

SHAPES = [(10, 10), (15, 7), (4, 12), (1, 1), (14, 15)]


def _inputs(h1, h2, B, N, seed, w_std=0.4):
    spec = O.FlowSpec(d=2, nb=1, hidden=[h1, h2], n_steps=N)
    params = O.init_params(spec, w_std, seed, torch.float32)
    g = torch.Generator().manual_seed(seed + 100)
    x_map = torch.randn(B, 2, generator=g).clamp(-3.0, 3.0)
    x_data = 1.3 * x_map.flip(-1) * torch.tensor([1.0, 0.6]) + torch.tensor([0.4, -0.2])
    return spec, params, x_data, x_map


def _cfg(h1, h2, N):
    from gpflow_b200 import FlowConfig
    return FlowConfig(d=2, nb=1, hidden=(h1, h2), n_steps=N)


def test_narrow_shapes_are_dispatched_to_the_padded_kernels():
    import gpflow_b200 as G
    from gpflow_b200 import _lib
    v = G.kernel_variant(_cfg(10, 10, 15))
    assert "sf2r" in v and "zero-padded" in v and "[10,10]" in v
    assert "sf2r_fwd" in G.kernel_variant(_cfg(10, 10, 15), op=_lib.OP_FORWARD)
    assert "sf2e" in G.kernel_variant(_cfg(15, 7, 15), euler_coef=0.3)
    assert "generic" in G.kernel_variant(_cfg(10, 10, 15), torch.float64)
    assert "generic" in G.kernel_variant(_cfg(10, 10, 15), force_generic=True)
    assert "zero-padded" not in G.kernel_variant(_cfg(16, 10, 15))      # wider than 15: tc / generic
    assert "zero-padded" not in G.kernel_variant(_cfg(15, 15, 15))


@pytest.mark.parametrize("h1,h2", SHAPES)
def test_train_step_vs_fp64_oracle_and_generic(h1, h2):
    import gpflow_b200 as G
    spec, params, x_data, x_map = _inputs(h1, h2, 300, 6, seed=10 * h1 + h2)
    r = O.train_step([p.double() for p in params], x_data.double(), x_map.double(), spec, need_grad_x=True)
    cfg = _cfg(h1, h2, 6)
    flat = O.flatten_params(params).cuda()
    assert flat.numel() == cfg.n_params
    res = G.fused_train_step(cfg, flat, x_data.cuda(), x_map.cuda(), want_y=True, want_grad_x=True)
    assert res["grads"].numel() == cfg.n_params
    assert rel(res["y"].cpu(), r["y"]) < 1e-5
    assert abs(float(res["scalars"][0]) - float(r["ot"])) < 1e-5 * abs(float(r["ot"]))
    assert rel(res["grads"].cpu(), r["grads"]) < 1e-5, rel(res["grads"].cpu(), r["grads"])
    assert rel(res["grad_x"].cpu(), r["grad_x"]) < 5e-5
    gen = G.fused_train_step(cfg, flat, x_data.cuda(), x_map.cuda(), want_y=True, force_generic=True)
    assert rel(res["grads"].cpu(), gen["grads"].cpu()) < 1e-5


def test_forward_reverse_and_logdet():
    import gpflow_b200 as G
    spec, params, x_data, x_map = _inputs(9, 13, 1000, 8, seed=3)
    cfg = _cfg(9, 13, 8)
    flat = O.flatten_params(params).cuda()
    a = G.flow_forward(cfg, flat, x_map.cuda(), wrap=True, want_logdet=True)
    b = G.flow_forward(cfg, flat, x_map.cuda(), wrap=True, want_logdet=True, force_generic=True)
    assert rel(a["y"].cpu(), b["y"].cpu()) < 2e-6
    # log-det = difference of the |x|^2 / 2 terms at both ends (network.py:170-189): compared on that scale, as in
    # tests/test_sf2_parity.py
    scale = float((b["y"].cpu().double() ** 2).sum(1).norm())
    assert float((a["logdet"].cpu().double() - b["logdet"].cpu().double()).norm()) < 5e-6 * scale
    back = G.flow_forward(cfg, flat, a["y"], wrap=True, reverse=True)
    back_gen = G.flow_forward(cfg, flat, a["y"], wrap=True, reverse=True, force_generic=True)
    assert rel(back["y"].cpu(), back_gen["y"].cpu()) < 2e-6
    assert rel(back["y"].cpu(), x_map) < 1e-3        # RK4 with 8 steps: the round trip carries the discretization error


def test_euler_penalized_step_on_a_narrow_net():
    import gpflow_b200 as G
    spec, params, x_data, x_map = _inputs(12, 6, 257, 5, seed=21, w_std=0.3)
    g = torch.Generator().manual_seed(5)
    zw = torch.randn(5, 2, 2, generator=g)
    r = O.train_step([p.double() for p in params], x_data.double(), x_map.double(), spec, euler_coef=0.3, zw=zw.double())
    r32 = O.train_step(params, x_data, x_map, spec, euler_coef=0.3, zw=zw)
    cfg = _cfg(12, 6, 5)
    res = G.fused_train_step(cfg, O.flatten_params(params).cuda(), x_data.cuda(), x_map.cuda(), euler_coef=0.3, zw=zw.cuda())
    sc = res["scalars"].cpu().double()
    assert abs(sc[0] - r["ot"]) < 1e-5 * abs(r["ot"])
    assert abs(sc[1] - r["reg"]) < 5e-3 * abs(r["reg"]) + 2 * abs(r32["reg"].double() - r["reg"])
    noise = rel(r32["grads"], r["grads"])       # fp32 noise of the finite-difference penalty, see tests/test_sf2e_parity.py
    assert rel(res["grads"].cpu(), r["grads"]) < 5e-3 + 2 * noise


def test_training_steps_with_the_fused_optimizer_keep_the_narrow_layout():
    """The module path: parameters of the narrow shape, gradients gathered back into it, Adam on the flat buffer."""
    import gpflow_b200 as G
    torch.manual_seed(0)
    flow = G.get_gp_flow(2, {"nb_neurons": [8, 11], "nb_div_blocks": 1}, 6).cuda()
    net = G.GPFlowNetwork(flow, map_module=None)
    opt = net.configure_fused_optimizer(lr=2e-3, weight_decay=1e-4)
    g = torch.Generator().manual_seed(1)
    x_map = torch.randn(4096, 2, generator=g).clamp(-3, 3).cuda()
    x_data = (1.2 * x_map.flip(-1) + 0.1)
    losses = []
    for _ in range(20):
        loss, reg = net.fused_training_step(x_data, x_map)
        opt.step()
        losses.append(float(loss))
    assert losses[-1] < losses[0]
    assert all(torch.isfinite(p).all() for p in flow.parameters())
