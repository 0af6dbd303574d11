"""GPU tests of the module-level surface the reference's callers use -- what `main.py --opt_type train_gp` actually runs
once INTEGRATION.md's hook is in place, as opposed to the tensor-level ops the other parity tests call:

  * `flow(x)` + `.backward()` through the autograd Function: gradients must land in `denses[i].weight.grad / .bias.grad`
    and in the input (src/gp_flows/network.py:86-114,227-236);
  * `Network.training_step`'s GP branch verbatim -- `loss_gp_flow(x)`; `reg = flow.regularization_coef()`;
    `zero_grad(); backward(); Adam.step()` -- replayed on the reference's own C1 / C2 training trajectories;
  * `save_trajectories=True` / `get_trajectories()` (res_net.py:248-283), `VelocityField.__call__` (velocity_field.py:118-126);
  * `Compose` (nb_blocks != 1, compose.py:5-69) forward, reverse, gradients against vectors from the unmodified reference;
  * the reference's own property test of this path, ported: the velocity is divergence-free
    (tests/test_velocity_field.py:41-96: d = 5, [5, 5, 5], nb_div_blocks in {1, 3}, weights N(0, 1)).
"""
import os

import numpy as np
import pytest
import torch

import gpflow_oracle as O
from golden_util import GOLDEN_DIR, load_golden, rel

pytestmark = pytest.mark.gpu


def _flow_from(c, euler=None, dtype=torch.float32):
    import gpflow_b200 as G
    spec = c["spec"]
    ed = {"use_euler": euler is not None, "case": "penalization", "coef_penalization": euler}
    flow = G.get_gp_flow(spec.d, {"nb_neurons": list(spec.hidden), "nb_div_blocks": spec.nb}, spec.n_steps, euler_dict=ed,
                         dtype=dtype).cuda()
    with torch.no_grad():
        for p, q in zip(flow.parameters(), c["params"]):
            p.copy_(q.to(dtype))
    return flow


@pytest.mark.parametrize("name", ["c1_d2_h15", "d5_nb3_L3_euler", "d4_nb4_L2"])
def test_autograd_through_the_module_matches_reference(name):
    """loss_gp_flow + backward through `_FlowFunction`; golden = the reference's own loss.backward()."""
    import gpflow_b200 as G
    c = load_golden(name, "f32")
    g = c["g"]
    flow = _flow_from(c, euler=c["euler"])
    net = G.GPFlowNetwork(flow, map_module=None)
    x_map = c["x_map"].cuda().requires_grad_("grad_x" in g)
    torch.manual_seed(1000 + {"c1_d2_h15": 1, "d5_nb3_L3_euler": 6, "d4_nb4_L2": 8}[name])   # seed_zw of make_golden.py
    y, logdet = net.gp_flow(x_map)
    loss = ((c["x_data"].cuda() - y) ** 2).sum(-1).mean()          # network.py:82
    reg = flow.regularization_coef()
    (loss + reg).backward()
    grads = torch.cat([p.grad.reshape(-1) for p in flow.parameters()]).cpu()
    tol = 5e-3 if c["euler"] is not None else 1e-5
    assert rel(y.detach().cpu(), g["y"]) < 1e-5
    assert rel(logdet.detach().cpu(), g["logdet"]) < 1e-4          # difference of two O(|y|^2) numbers
    assert abs(float(loss) - float(g["ot"])) < 1e-5 * abs(float(g["ot"]))
    assert rel(grads, g["grads"]) < tol, rel(grads, g["grads"])
    for dense in flow.v.v.denses:
        assert dense.weight.grad is not None and dense.bias.grad is not None
    if "grad_x" in g:
        assert rel(x_map.grad.cpu(), g["grad_x"]) < max(tol, 5e-5), rel(x_map.grad.cpu(), g["grad_x"])
    assert abs(float(flow.lagrangian) - float(g["lagrangian"])) < 1e-5 * abs(float(g["lagrangian"])) + 1e-12


@pytest.mark.parametrize("name", ["train_c1_eight_gaussians", "train_c2_moons_euler"])
def test_training_step_as_the_reference_writes_it(name):
    """network.py:86-114 verbatim on the module mirror (autograd path, torch.optim.Adam), against the K optimiser steps the
    unmodified reference ran (oracle/make_golden_train.py)."""
    import gpflow_b200 as G
    z = np.load(os.path.join(GOLDEN_DIR, name + "_f32.npz"))
    g = {k: z[k] for k in z.files}
    euler = None if float(g["euler_coef"]) < 0 else float(g["euler_coef"])
    spec = O.FlowSpec(d=2, nb=1, hidden=[15, 15], n_steps=15)
    flow = G.get_gp_flow(2, {"nb_neurons": [15, 15]}, 15,
                         euler_dict={"use_euler": euler is not None, "case": "penalization", "coef_penalization": euler}).cuda()
    s, off, flat0 = spec.layer_sizes(), 0, torch.from_numpy(g["params0"])
    with torch.no_grad():
        for p in flow.parameters():
            p.copy_(flat0[off:off + p.numel()].view_as(p)); off += p.numel()
    net = G.GPFlowNetwork(flow, map_module=None)
    opt = torch.optim.Adam(flow.parameters(), lr=2e-3, betas=(0.9, 0.999), weight_decay=1e-4)   # network.py:312-317
    for k in range(int(g["K"])):
        torch.manual_seed(5000 + k)
        x, x_map = torch.from_numpy(g["x"][k]).cuda(), torch.from_numpy(g["x_map"][k]).cuda()
        loss_gp = net.loss_gp_flow(x, x_map)
        reg = flow.regularization_coef()
        total = loss_gp + reg
        opt.zero_grad()
        total.backward()
        opt.step()
        assert abs(float(loss_gp) - g["ot"][k]) < 1e-5 * abs(g["ot"][k]), (k, float(loss_gp), g["ot"][k])
        if euler is None:
            assert float(reg) == 0.0
    flat = torch.cat([p.detach().reshape(-1) for p in flow.parameters()]).cpu()
    assert rel(flat, g["paramsK"]) < 2e-4, rel(flat, g["paramsK"])


def test_save_trajectories_and_velocity_call():
    import gpflow_b200 as G
    c = load_golden("c1_d2_h15", "f32")
    g = c["g"]
    flow = _flow_from(c)
    g0 = torch.erf(c["x_map"] / np.sqrt(2.0)).cuda()
    with torch.no_grad():
        gT = flow(g0, save_trajectories=True)
    assert rel(gT.cpu(), g["g_T"]) < 2e-6
    xs, vs = flow.get_trajectories()
    # process_ode_solution (res_net.py:248-272): x_0, then every step with i % max(nb_layers // 40, 1) == 0 -> all 15
    assert xs.shape == (16, 64, 2) and vs.shape == (16, 64, 2)
    assert rel(xs, g["traj"]) < 2e-6, rel(xs, g["traj"])
    # v_trajectories[k] = v(x_0, t_k) (the reference evaluates the velocity at the START point for the successive times)
    p64 = [p.double() for p in c["params"]]
    eps = O.eps_of(torch.float32)
    t = torch.tensor(0.0)
    dt = torch.tensor(1.0 / 15)
    for k in range(16):
        v_ref = O.velocity(p64, g0.cpu().double(), t.double(), 1, eps)
        assert rel(vs[k], v_ref) < 1e-5, (k, rel(vs[k], v_ref))
        t = t + dt
    # VelocityField.__call__ on its own, fp32 and fp64
    v32 = flow.v(g0, 0.4)
    assert rel(v32.cpu(), O.velocity(p64, g0.cpu().double(), torch.tensor(0.4, dtype=torch.float64), 1, eps)) < 1e-5
    flow64 = _flow_from(c, dtype=torch.float64)
    p64b = [p.detach().cpu() for p in flow64.parameters()]
    v64 = flow64.v(g0.double(), 0.4)
    assert rel(v64.cpu(), O.velocity(p64b, g0.cpu().double(), torch.tensor(0.4, dtype=torch.float64), 1,
                                      O.eps_of(torch.float64))) < 1e-10
    # reverse trajectories: same count, start = the given points
    with torch.no_grad():
        flow(gT, reverse=True, save_trajectories=True)
    xr, _ = flow.get_trajectories()
    assert xr.shape == (16, 64, 2) and rel(xr[0], gT.cpu()) == 0.0 and rel(xr[-1], g0.cpu()) < 1e-4


@pytest.mark.parametrize("name", ["compose_d2_h15_b2", "compose_d3_nb2_b3"])
def test_compose_blocks_match_reference(name):
    """nb_blocks != 1: forward, log-det, OT cost, per-block gradients, input gradient, reverse, Lagrangian sum."""
    import gpflow_b200 as G
    from gpflow_b200.res_net import Compose
    z = np.load(os.path.join(GOLDEN_DIR, name + "_f32.npz"))
    g = {k: z[k] for k in z.files}
    d, nb, H, N, nblk = int(g["d"]), int(g["nb"]), [int(h) for h in g["hidden"]], int(g["N"]), int(g["blocks"])
    flow = G.get_gp_flow(d, {"nb_neurons": H, "nb_div_blocks": nb}, N, nb_blocks=nblk).cuda()
    assert isinstance(flow, Compose) and len(flow.layers) == nblk
    with torch.no_grad():
        for bi, layer in enumerate(flow.layers):
            flat, off = torch.from_numpy(g["params"][bi]), 0
            for p in layer.parameters():
                p.copy_(flat[off:off + p.numel()].view_as(p)); off += p.numel()
    net = G.GPFlowNetwork(flow, map_module=None)
    x_map = torch.from_numpy(g["x_map"]).cuda().requires_grad_(True)
    x_data = torch.from_numpy(g["x_data"]).cuda()
    y, logdet = net.gp_flow(x_map)
    ot = ((x_data - y) ** 2).sum(-1).mean()
    (ot + flow.regularization_coef()).backward()
    assert rel(y.detach().cpu(), g["y"]) < 1e-5
    assert abs(float(ot) - float(g["ot"])) < 1e-5 * float(g["ot"])
    assert abs(float(flow.lagrangian) - float(g["lagrangian"])) < 1e-5 * float(g["lagrangian"])
    for bi, layer in enumerate(flow.layers):
        gr = torch.cat([p.grad.reshape(-1) for p in layer.parameters()]).cpu()
        assert rel(gr, g["grads"][bi]) < 1e-5, (bi, rel(gr, g["grads"][bi]))
    assert rel(x_map.grad.cpu(), g["grad_x"]) < 5e-5
    with torch.no_grad():
        y_rev = net.gp_flow(x_data, reverse=True)
    assert rel(y_rev.cpu(), g["y_rev"]) < 1e-5
    # the fused-step entry point must also work for a composed flow (autograd chain), same numbers
    for p in flow.parameters():
        p.grad = None
    loss, reg = net.fused_training_step(x_data, x_map.detach())
    assert abs(float(loss) - float(g["ot"])) < 1e-5 * float(g["ot"]) and float(reg) == 0.0
    gr0 = torch.cat([p.grad.reshape(-1) for p in flow.layers[0].parameters()]).cpu()
    assert rel(gr0, g["grads"][0]) < 1e-5


@pytest.mark.parametrize("nb", [1, 3])
def test_velocity_is_divergence_free(nb):
    """Port of /root/reference/tests/test_velocity_field.py:41-96 to the CUDA velocity: dim 5, [5, 5, 5], weights N(0, 1)
    (`init_v`, :14-23).  The reference bounds the autograd divergence by 5e-5 in fp32; here the divergence of the fp64
    kernel is taken by central finite differences (SURVEY.md 8c asks < 1e-8), and the fp32 kernel is held to the fp64
    one.  Also: v . n = 0 on the walls x_i^2 = 1 - eps (the velocity is wall-tangent)."""
    import gpflow_b200 as G
    d, H = 5, [5, 5, 5]
    torch.manual_seed(123)
    flow = G.get_gp_flow(d, {"nb_neurons": H, "nb_div_blocks": nb}, 10, dtype=torch.float64).cuda()
    with torch.no_grad():
        for p in flow.parameters():
            p.copy_(torch.randn(p.shape, dtype=torch.float64))
    x = (torch.rand(50, d, dtype=torch.float64) * 1.6 - 0.8).cuda()
    t = 0.37
    h = 1e-4          # 4th-order central differences: truncation h^4 |v^(5)| / 30 and rounding eps |v| / h are both ~1e-10 here
    div = torch.zeros(50, dtype=torch.float64, device="cuda")
    for j in range(d):
        e = torch.zeros(d, dtype=torch.float64, device="cuda"); e[j] = h
        div += (8 * (flow.v(x + e, t)[:, j] - flow.v(x - e, t)[:, j]) - (flow.v(x + 2 * e, t)[:, j] - flow.v(x - 2 * e, t)[:, j])) / (12 * h)
    vmag = float(flow.v(x, t).norm(dim=1).max())
    assert vmag > 1.0                                        # a velocity that is not trivially small
    assert float(div.abs().max()) < 1e-8, (float(div.abs().max()), vmag)      # fp64 FD divergence < 1e-8
    # fp32 kernel against the fp64 one
    flow32 = G.get_gp_flow(d, {"nb_neurons": H, "nb_div_blocks": nb}, 10).cuda()
    with torch.no_grad():
        for p, q in zip(flow32.parameters(), flow.parameters()):
            p.copy_(q.float())
    flow_rt = G.get_gp_flow(d, {"nb_neurons": H, "nb_div_blocks": nb}, 10, dtype=torch.float64).cuda()
    with torch.no_grad():
        for p, q in zip(flow_rt.parameters(), flow32.parameters()):
            p.copy_(q.double())
    xf = x.float()
    assert rel(flow32.v(xf, t).cpu(), flow_rt.v(xf.double(), t).cpu()) < 1e-5
    # wall tangency: put coordinate 2 on the wall
    xw = x.clone()
    xw[:, 2] = float(np.sqrt(flow.v.bound))
    assert float(flow.v(xw, t)[:, 2].abs().max()) < 1e-12 * max(vmag, 1.0)


def test_compute_ot_costs_matches_reference_report():
    """`compute_ot_costs` (src/plots/outputs.py:7-248) on the pretrained C1 map (BNAF mirror) + a seeded GP flow: the numbers
    of the report the unmodified reference printed (tests/golden/ot_costs_c1_f32.npz, oracle/make_golden_ot_costs.py) --
    NLL of the map and of the OT map, both OT costs, the Lagrangian on the 100 x 100 mesh -- and the report text itself."""
    import gpflow_b200 as G
    from gpflow_b200.ot_costs import compute_ot_costs
    z = np.load(os.path.join(GOLDEN_DIR, "ot_costs_c1_f32.npz"))
    b = np.load(os.path.join(GOLDEN_DIR, "bnaf_eight_gaussians_f32.npz"))
    bmap = G.BNAFMap(int(b["dim"]), int(b["n_neurons"]), int(b["n_layers"]), torch.from_numpy(b["params"])).cuda()
    flow = G.get_gp_flow(2, {"nb_neurons": [15, 15]}, 15).cuda()
    flat, off = torch.from_numpy(z["flow_params"]), 0
    with torch.no_grad():
        for p in flow.parameters():
            p.copy_(flat[off:off + p.numel()].view_as(p)); off += p.numel()
    net = G.GPFlowNetwork(flow, map_module=bmap)
    x_test = torch.from_numpy(z["x_test"]).cuda()
    res, text = compute_ot_costs(net, x_test, batch_size=1024)
    for k in ("nll_map", "cost_map", "nll_ot", "cost_gp_map", "lagrangian"):
        assert abs(res[k] - float(z[k])) < 2e-5 * abs(float(z[k])), (k, res[k], float(z[k]))
    # Gaussian-preserving: the OT map has the same likelihood as the map it composes with (to fp32 accuracy)
    assert abs(res["nll_ot"] - res["nll_map"]) < 1e-4
    assert text.strip() == str(z["report"]).strip()
    with torch.no_grad():
        y_ot, ld = net.ot_map(x_test, return_log_det=True)
    assert rel(y_ot.cpu(), z["y_ot"]) < 1e-5 and rel(ld.cpu(), z["log_det_ot"]) < 1e-4


def test_compute_ot_costs_inverse_consistency_with_ffjord_map():
    """The block the reference runs when the map has a `backward` (outputs.py:101-137): finv_f errors and backward OT
    costs, here with the FFJORD mirror of the pretrained dsprites map (d = 10: Lagrangian on uniform cube samples)."""
    import gpflow_b200 as G
    from gpflow_b200.ot_costs import compute_ot_costs
    z = np.load(os.path.join(GOLDEN_DIR, "ffjord_dsprites_f32.npz"))
    m = G.FFJORDMap(int(z["d"]), [int(h) for h in z["hidden"]], int(z["n_blocks"]), torch.from_numpy(z["params"]),
                    [float(t) for t in z["t_end"]], 64).cuda()
    torch.manual_seed(3)
    flow = G.get_gp_flow(10, {"nb_neurons": [50, 50, 50], "nb_div_blocks": 9}, 15).cuda()
    with torch.no_grad():
        for p in flow.parameters():
            p.normal_(0.0, 0.03)
    net = G.GPFlowNetwork(flow, map_module=m)
    x_gauss = torch.randn(600, 10, generator=torch.Generator().manual_seed(4)).cuda()
    x_test = m.backward(torch.randn(600, 10, generator=torch.Generator().manual_seed(5)).cuda())
    res, text = compute_ot_costs(net, x_test, batch_size=256, x_gauss=x_gauss)
    assert res["finv_f_err"] < 1e-4 and res["finv_f_gp_err"] < 2e-3        # RK4 round trips (64 / 15 steps)
    assert res["cost_back"] > 0 and res["cost_gp_back"] > 0 and res["lagrangian"] > 0
    assert abs(res["nll_ot"] - res["nll_map"]) < 5e-3
    assert "finv_f error GP map" in text and "OT cost GP backward random data" in text
