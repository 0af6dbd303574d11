"""Errors of the C4-shape fused train step against the fp64 oracle on identical inputs (B = 4099, N = 15, weights
N(0, 0.3^2), x_map clamped to 3 sigma like tests/test_sf2_parity.py), for the kernel the current build / environment
selects (GPFLOW_SF2_FORM=0: first formulation; force_v1: the register-resident kernel).  Prints one line per variant."""
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "oracle"))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import gpflow_oracle as O  # noqa: E402
from golden_util import rel  # noqa: E402
import gpflow_b200 as G  # noqa: E402

SPEC = O.FlowSpec(d=2, nb=1, hidden=[15, 15], n_steps=15)
cfg = G.FlowConfig(d=2, nb=1, hidden=(15, 15), n_steps=15)
for seed, B, w_std in ((4099, 4099, 0.3), (7, 2048, 0.6), (11, 2048, 0.1)):
    params = O.init_params(SPEC, w_std, seed, torch.float32)
    g = torch.Generator().manual_seed(seed + 100)
    x_map = torch.randn(B, 2, generator=g).clamp(-3, 3)
    x_data = 1.3 * x_map.flip(-1) * torch.tensor([1.0, 0.6]) + torch.tensor([0.4, -0.2])
    r = O.train_step([p.double() for p in params], x_data.double(), x_map.double(), SPEC, need_grad_x=True)
    flat = O.flatten_params(params).cuda()
    for label, kw in (("default", {}), ("v1", {"force_v1": True})):
        res = G.fused_train_step(cfg, flat, x_data.cuda(), x_map.cuda(), want_y=True, want_grad_x=True, **kw)
        sc = res["scalars"].cpu().double()
        print("w_std=%.1f B=%d %-8s %-55s y %.2e  ot %.2e  grads %.2e  grad_x %.2e" % (
            w_std, B, label, G.kernel_variant(cfg, **kw), rel(res["y"].cpu(), r["y"]), abs(float(sc[0] - r["ot"]) / float(r["ot"])),
            rel(res["grads"].cpu(), r["grads"]), rel(res["grad_x"].cpu(), r["grad_x"])))
