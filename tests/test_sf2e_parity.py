"""GPU parity tests of `gpflow_sf2e_kernel`: the 2-D train step / backward op WITH the Euler penalization
(res_net.py:291-302,319-358) on the sf2r structure, the dual number held in an fp32x2 register pair (gpflow_sf2e*.cuh).

The penalty is a finite difference with delta = 2 sqrt(eps) = 6.9e-4 evaluated in fp32: its value and gradient carry
~1e-3 of rounding noise whatever the implementation (the reference's own fp32 run is that far from its fp64 run, SURVEY.md
section 6), so the bar against the fp64 oracle is the 5e-3 of tests/test_cuda_parity.py.  The tight checks are structural:
  * the OT part of the kernel (penalty cotangent 0) against the plain sf2r kernel, 1e-5;
  * linearity of the gradient in the cotangent of the regularization value (backward op);
  * the independent v1 dual kernel on ragged multi-tile batches;
  * additivity over batch shards (what the data-parallel all-reduce relies on), determinism, out-of-range lanes.
"""
import numpy as np
import pytest
import torch

import gpflow_oracle as O
from golden_util import rel

pytestmark = pytest.mark.gpu

SPEC = O.FlowSpec(d=2, nb=1, hidden=[15, 15], n_steps=15)
EULER = 0.3          # large enough for the penalty gradient to matter next to the OT gradient (as in the goldens)


def _cfg():
    from gpflow_b200 import FlowConfig
    return FlowConfig(d=2, nb=1, hidden=(15, 15), n_steps=15)


def _inputs(B, seed, w_std=0.3):
    params = O.init_params(SPEC, w_std, seed, torch.float32)
    g = torch.Generator().manual_seed(seed + 100)
    x_map = torch.randn(B, 2, generator=g).clamp(-3.0, 3.0)
    x_data = 1.3 * x_map.flip(-1) * torch.tensor([1.0, 0.6]) + torch.tensor([0.4, -0.2])
    zw = torch.randn(SPEC.n_steps, 2, 2, generator=g)
    return params, x_data, x_map, zw


def test_sf2e_is_dispatched():
    import gpflow_b200 as G
    from gpflow_b200 import _lib
    assert "sf2e" in G.kernel_variant(_cfg(), euler_coef=EULER)
    assert "sf2e" in G.kernel_variant(_cfg(), euler_coef=EULER, op=_lib.OP_BACKWARD)
    assert "small" in G.kernel_variant(_cfg(), euler_coef=EULER, force_v1=True)
    assert "small" in G.kernel_variant(_cfg(), euler_coef=EULER, op=_lib.OP_FORWARD)      # forward op: v1 kernel
    assert "sf2r" in G.kernel_variant(_cfg())


@pytest.mark.parametrize("B", [1, 33, 257])
def test_sf2e_train_step_vs_fp64_oracle(B):
    import gpflow_b200 as G
    params, x_data, x_map, zw = _inputs(B, seed=40 + B)
    r = O.train_step([p.double() for p in params], x_data.double(), x_map.double(), SPEC, euler_coef=EULER, zw=zw.double(),
                     need_grad_x=True)
    # the same formulas in torch fp32 on the CPU: how far fp32 itself is from fp64 on this penalty-dominated gradient (the
    # difference quotient divides fp32 rounding by delta = 6.9e-4)
    r32 = O.train_step(params, x_data, x_map, SPEC, euler_coef=EULER, zw=zw, need_grad_x=True)
    res = G.fused_train_step(_cfg(), O.flatten_params(params).cuda(), x_data.cuda(), x_map.cuda(), euler_coef=EULER,
                             zw=zw.cuda(), want_y=True, want_grad_x=True)
    sc = res["scalars"].cpu().double()
    assert rel(res["y"].cpu(), r["y"]) < 1e-5
    assert abs(sc[0] - r["ot"]) < 1e-5 * abs(r["ot"])
    assert abs(sc[1] - r["reg"]) < 5e-3 * abs(r["reg"]) + 2 * abs(r32["reg"].double() - r["reg"]), (float(sc[1]), float(r["reg"]))
    assert float(sc[1]) > 0.0
    noise_g, noise_x = rel(r32["grads"], r["grads"]), rel(r32["grad_x"], r["grad_x"])
    assert rel(res["grads"].cpu(), r["grads"]) < 5e-3 + 2 * noise_g, (rel(res["grads"].cpu(), r["grads"]), noise_g)
    assert rel(res["grad_x"].cpu(), r["grad_x"]) < 5e-3 + 2 * noise_x, (rel(res["grad_x"].cpu(), r["grad_x"]), noise_x)


@pytest.mark.parametrize("B", [31, 4099, 148 * 8 * 32 + 5])
def test_sf2e_matches_v1_dual_kernel(B):
    """Ragged, multi-tile batches (the last one gives every resident warp a tile and one warp a second, ragged one)."""
    import gpflow_b200 as G
    params, x_data, x_map, zw = _inputs(B, seed=7 + B % 1000)
    flat = O.flatten_params(params).cuda()
    kw = dict(euler_coef=EULER, zw=zw.cuda(), want_y=True, want_grad_x=True)
    a = G.fused_train_step(_cfg(), flat, x_data.cuda(), x_map.cuda(), **kw)
    b = G.fused_train_step(_cfg(), flat, x_data.cuda(), x_map.cuda(), force_v1=True, **kw)
    assert rel(a["y"].cpu(), b["y"].cpu()) < 2e-6
    assert abs(float(a["scalars"][0]) - float(b["scalars"][0])) < 2e-6 * abs(float(b["scalars"][0]))
    assert abs(float(a["scalars"][1]) - float(b["scalars"][1])) < 5e-3 * abs(float(b["scalars"][1]))
    assert rel(a["grads"].cpu(), b["grads"].cpu()) < 5e-3, rel(a["grads"].cpu(), b["grads"].cpu())
    assert rel(a["grad_x"].cpu(), b["grad_x"].cpu()) < 5e-3
    a2 = G.fused_train_step(_cfg(), flat, x_data.cuda(), x_map.cuda(), **kw)          # fixed-order reductions
    assert torch.equal(a["buf"], a2["buf"]) and torch.equal(a["grad_x"], a2["grad_x"])


def test_sf2e_backward_op_is_linear_in_the_regularization_cotangent():
    """grads(c) = grads_OT + c * grads_pen for the cotangent c of the regularization value (network.py:104-111 adds
    reg to the loss with weight 1): c = 0 must reproduce the PLAIN kernel's gradient (the OT part of sf2e, tight),
    and the c-dependence must be linear; grads_pen against the fp64 oracle at the finite-difference bar."""
    import gpflow_b200 as G
    B = 1500
    params, x_data, x_map, zw = _inputs(B, seed=3)
    flat = O.flatten_params(params).cuda()
    xm = x_map.cuda()
    gy = (0.01 * torch.randn(B, 2, generator=torch.Generator().manual_seed(5))).cuda()

    def bwd(c):
        return G.flow_backward(_cfg(), flat, xm, gy, wrap=True, training=True, euler_coef=EULER, zw=zw.cuda(),
                               reg_cotangent=torch.tensor(float(c), device="cuda"), loss_scale=1.0 / B)

    g0, x0 = bwd(0.0)
    g1, x1 = bwd(1.0)
    g3, x3 = bwd(3.0)
    gp, xp = G.flow_backward(_cfg(), flat, xm, gy, wrap=True, training=True, loss_scale=1.0 / B)   # sf2r, no penalty
    assert rel(g0.cpu(), gp.cpu()) < 1e-5, rel(g0.cpu(), gp.cpu())
    assert rel(x0.cpu(), xp.cpu()) < 1e-5
    pen1, pen3 = (g1 - g0).cpu().double(), (g3 - g0).cpu().double()
    assert rel(pen3, 3.0 * pen1) < 1e-4, rel(pen3, 3.0 * pen1)
    assert rel((x3 - x0).cpu(), 3.0 * (x1 - x0).cpu()) < 1e-3      # differences of fp32 numbers
    # the penalty gradient itself, against the fp64 oracle: d(reg)/d(params) = grads(euler) - grads(no euler) at OT seed 0
    ps = [p.double().requires_grad_(True) for p in params]
    _, _, res = O.gp_flow(ps, x_map.double(), SPEC, euler_coef=EULER, zw=zw.double(), with_lagrangian=False)
    res.reg.backward()
    ref = torch.cat([p.grad.reshape(-1) for p in ps])
    ps32 = [p.clone().requires_grad_(True) for p in params]
    _, _, res32 = O.gp_flow(ps32, x_map, SPEC, euler_coef=EULER, zw=zw, with_lagrangian=False)
    res32.reg.backward()
    noise = rel(torch.cat([p.grad.reshape(-1) for p in ps32]), ref)      # fp32 torch against fp64, same formulas
    assert rel(pen1, ref) < 5e-3 + 2 * noise, (rel(pen1, ref), noise)


def test_sf2e_buffer_is_additive_over_batch_shards():
    import gpflow_b200 as G
    B = 3000
    params, x_data, x_map, zw = _inputs(B, seed=11)
    flat = O.flatten_params(params).cuda()
    kw = dict(euler_coef=EULER, zw=zw.cuda())
    full = G.fused_train_step(_cfg(), flat, x_data.cuda(), x_map.cuda(), **kw)["buf"].cpu().double()
    cuts = [0, 700, 731, 2000, B]
    acc = torch.zeros_like(full)
    for lo, hi in zip(cuts[:-1], cuts[1:]):
        acc += G.fused_train_step(_cfg(), flat, x_data[lo:hi].cuda(), x_map[lo:hi].cuda(), loss_scale=1.0 / B,
                                  **kw)["buf"].cpu().double()
    assert rel(acc, full) < 5e-5, rel(acc, full)


def test_sf2e_small_coefficient_and_reference_init():
    """The reference's setting: weights N(0, 1e-3^2), coefficient 5e-4 (default_params/toy_params.py): the penalty is tiny
    there; value and total gradient against the fp64 oracle."""
    import gpflow_b200 as G
    B = 512
    params = O.init_params(SPEC, 1e-3, 21, torch.float32)
    g = torch.Generator().manual_seed(77)
    x_map = torch.randn(B, 2, generator=g).clamp(-3.0, 3.0)
    x_data = x_map + 0.1 * torch.randn(B, 2, generator=g)
    zw = torch.randn(SPEC.n_steps, 2, 2, generator=g)
    r = O.train_step([p.double() for p in params], x_data.double(), x_map.double(), SPEC, euler_coef=5e-4, zw=zw.double())
    res = G.fused_train_step(_cfg(), O.flatten_params(params).cuda(), x_data.cuda(), x_map.cuda(), euler_coef=5e-4,
                             zw=zw.cuda(), want_y=True)
    assert "sf2e" in G.kernel_variant(_cfg(), euler_coef=5e-4)
    assert rel(res["y"].cpu(), r["y"]) < 1e-5
    assert abs(float(res["scalars"][0]) - float(r["ot"])) < 1e-5 * abs(float(r["ot"]))
    assert rel(res["grads"].cpu(), r["grads"]) < 1e-5 + 5e-3 * float(r["reg"]) / max(float(r["ot"]), 1e-30)
    assert np.isfinite(float(res["scalars"][1]))
