"""Training-trajectory parity on the reference's real C1 / C2 pipeline pieces: the K optimiser steps the
UNMODIFIED reference ran on CPU (toy data generator, pretrained BNAF map, its own get_gp_flow / loss_gp_flow /
regularization_coef / Adam; oracle/make_golden_train.py) are replayed
  * by the CPU oracle (pins the oracle on a multi-step run; not gpu), and
  * through the CUDA path: gpflow_b200.get_gp_flow + GPFlowNetwork.fused_training_step + torch Adam (gpu),
on the identical batches, map outputs and Euler probes.  Compared: the per-step OT cost, the regulariser, the
parameters after K steps.  Both sides are fp32 and Adam divides by sqrt(v): 1e-5 on the losses, 2e-4 on the weights."""
import os

import numpy as np
import pytest
import torch

import gpflow_oracle as O
from golden_util import GOLDEN_DIR, rel

CASES = ["train_c1_eight_gaussians", "train_c2_moons_euler"]
SPEC = O.FlowSpec(d=2, nb=1, hidden=[15, 15], n_steps=15)


def _load(name):
    z = np.load(os.path.join(GOLDEN_DIR, name + "_f32.npz"))
    g = {k: z[k] for k in z.files}
    g["euler"] = None if float(g["euler_coef"]) < 0 else float(g["euler_coef"])
    return g


def _unflatten(flat):
    ps, off, s = [], 0, SPEC.layer_sizes()
    for i in range(len(s) - 1):
        ps.append(flat[off: off + s[i] * s[i + 1]].reshape(s[i + 1], s[i]).clone()); off += s[i] * s[i + 1]
        ps.append(flat[off: off + s[i + 1]].clone()); off += s[i + 1]
    return ps


@pytest.mark.parametrize("name", CASES)
def test_oracle_replays_reference_training(name):
    g = _load(name)
    params = [p.requires_grad_(True) for p in _unflatten(torch.from_numpy(g["params0"]))]
    opt = torch.optim.Adam(params, lr=2e-3, betas=(0.9, 0.999), weight_decay=1e-4)
    K = int(g["K"]) if g["euler"] is None else 3          # the Euler oracle is slow on CPU: 3 steps are enough here
    for k in range(K):
        r = O.train_step([p.detach() for p in params], torch.from_numpy(g["x"][k]), torch.from_numpy(g["x_map"][k]),
                         SPEC, euler_coef=g["euler"], zw=torch.from_numpy(g["zw"][k]))
        assert abs(float(r["ot"]) - g["ot"][k]) < 1e-5 * abs(g["ot"][k]), (k, float(r["ot"]), g["ot"][k])
        off = 0
        for p in params:
            p.grad = r["grads"][off: off + p.numel()].reshape(p.shape).clone(); off += p.numel()
        opt.step()
    if K == int(g["K"]):
        flat = torch.cat([p.detach().reshape(-1) for p in params])
        assert rel(flat, g["paramsK"]) < 2e-4, rel(flat, g["paramsK"])


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_cuda_path_replays_reference_training(name):
    import gpflow_b200 as G
    g = _load(name)
    euler = g["euler"]
    flow = G.get_gp_flow(2, {"nb_neurons": [15, 15]}, 15,
                         euler_dict={"use_euler": euler is not None, "case": "penalization",
                                     "coef_penalization": euler}).cuda()
    with torch.no_grad():
        for p, q in zip(flow.parameters(), _unflatten(torch.from_numpy(g["params0"]))):
            p.copy_(q)
    net = G.GPFlowNetwork(flow, map_module=None)
    opt = torch.optim.Adam(flow.parameters(), lr=2e-3, betas=(0.9, 0.999), weight_decay=1e-4)   # network.py:312-317
    expect = "sf2" if euler is None else "euler"
    assert expect in G.kernel_variant(flow.config(), euler_coef=euler)
    for k in range(int(g["K"])):
        torch.manual_seed(5000 + k)                 # the wrapper draws z, w with the reference's RNG calls
        if euler is not None:
            state = torch.get_rng_state()
            assert torch.equal(flow.draw_probes(15, torch.zeros(1)), torch.from_numpy(g["zw"][k]))
            torch.set_rng_state(state)
        loss, reg = net.fused_training_step(torch.from_numpy(g["x"][k]).cuda(), torch.from_numpy(g["x_map"][k]).cuda())
        assert abs(float(loss) - g["ot"][k]) < 1e-5 * abs(g["ot"][k]), (k, float(loss), g["ot"][k])
        if euler is not None:
            # fp32 finite difference with delta = 7e-4 of a ~1e-3-amplitude field: the reference's own value is rounding
            # noise dominated (1e-16 .. 1e-8 here); bound it by the size of the OT cost instead of relatively
            assert abs(float(reg) - g["reg"][k]) < 1e-6 * abs(g["ot"][k])
        else:
            assert float(reg) == 0.0
        opt.step()
    flat = torch.cat([p.detach().reshape(-1) for p in flow.parameters()]).cpu()
    assert rel(flat, g["paramsK"]) < 2e-4, rel(flat, g["paramsK"])


# ---- the first 100 optimiser steps of C1 - C3 ------------------------------------------------------------
# Fixtures: oracle/make_golden_train.py (train100_*: the unmodified reference, batch 256 / 256 / 64; C3 on a synthetic map
# output because its latent pipeline is out of scope).  Over 100 Adam steps the fp32 rounding differences of the first steps
# are carried along by the optimiser (Adam divides by sqrt(v), and at the reference's near-identity initialisation the
# gradients are ~1e-6): measured on a B200 (tools/train100_dev.py) the per-step OT cost stays within 2.4e-5 of the
# reference's and the parameters after 100 steps within 4.4e-4 (C1), 3.1e-6 (C2 + Euler), 2.5e-5 (C3), while they move by
# 100 % of their norm.  Already at step 0 (identical parameters) the 256-sample OT cost is 1.9e-5 from the reference's: these
# small batches contain tail samples whose erfinv amplifies fp32 rounding (tests/test_sf2_parity.py).
# Bars: 5e-5 on every per-step OT cost, 2e-3 on the weights after 100 steps.
CASES100 = ["train100_c1_eight_gaussians", "train100_c2_moons_euler", "train100_c3_latent_synthetic"]


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES100)
def test_cuda_path_replays_100_reference_steps(name):
    import gpflow_b200 as G
    g = _load(name)
    euler = g["euler"]
    d, nb, hidden = int(g["d"]), int(g["nb"]), [int(h) for h in g["hidden"]]
    flow = G.get_gp_flow(d, {"nb_neurons": hidden, "nb_div_blocks": nb}, 15,
                         euler_dict={"use_euler": euler is not None, "case": "penalization",
                                     "coef_penalization": euler}).cuda()
    off, flat0 = 0, torch.from_numpy(g["params0"])
    with torch.no_grad():
        for p in flow.parameters():
            p.copy_(flat0[off:off + p.numel()].view_as(p)); off += p.numel()
    net = G.GPFlowNetwork(flow, map_module=None)
    opt = torch.optim.Adam(flow.parameters(), lr=2e-3, betas=(0.9, 0.999), weight_decay=1e-4)   # network.py:312-317
    K = int(g["K"])
    assert K == 100
    for k in range(K):
        torch.manual_seed(5000 + k)                 # the wrapper draws z, w with the reference's RNG calls
        loss, reg = net.fused_training_step(torch.from_numpy(g["x"][k]).cuda(), torch.from_numpy(g["x_map"][k]).cuda())
        assert abs(float(loss) - g["ot"][k]) < 5e-5 * abs(g["ot"][k]), (k, float(loss), g["ot"][k])
        if euler is not None:
            assert abs(float(reg) - g["reg"][k]) < 1e-6 * abs(g["ot"][k])
        else:
            assert float(reg) == 0.0
        opt.step()
    flat = torch.cat([p.detach().reshape(-1) for p in flow.parameters()]).cpu()
    assert rel(flat, g["paramsK"]) < 2e-3, rel(flat, g["paramsK"])
    assert rel(flat, flat0) > 0.5                     # the parameters really moved
