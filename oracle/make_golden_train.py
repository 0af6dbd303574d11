"""TEST INFRASTRUCTURE ONLY -- golden TRAINING TRAJECTORIES produced by the UNMODIFIED reference (needs /root/reference).

    python oracle/make_golden_train.py

BASELINE.json configs C1 / C2 on the reference's real pipeline pieces (SURVEY.md 8d "C1-C3 inputs"):
  * data        `inf_train_gen(<eight_gaussians|moons>)`            src/datasets/distributions_toy.py:12-85
  * frozen map  pretrained_models/<data>/bnaf/torch_nf_model.pt      flow_models/bnaf, `map(x)[0]` (network.py:76-78)
  * GP flow     `get_gp_flow(2, {"nb_neurons": [15, 15]}, 15, ...)`  res_net.py:14-67, reference init after manual_seed(0)
  * one step    `Network.loss_gp_flow` + `regularization_coef()` + zero_grad / backward / Adam.step
                (network.py:70-84,104-111), Adam(lr 2e-3, betas (0.9, 0.999), weight_decay 1e-4) (network.py:312-317,
                toy_params.py:94-104, src/data.py:89)
K optimiser steps at batch 1024 are run with the reference's own classes on CPU; stored per case (fp32):
the K batches `x`, the map outputs `x_map`, the Euler probes the reference drew (C2), the initial flat parameters,
the per-step OT cost / regulariser, and the flat parameters after the K steps.  `tests/test_train_trajectory.py`
replays the same K steps through the CUDA path.
"""
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")
REF = "/root/reference"

CASES = [
    dict(name="train_c1_eight_gaussians", data="eight_gaussians", K=12, B=1024, euler=None),
    dict(name="train_c2_moons_euler", data="moons", K=8, B=1024, euler=5e-4),
    # "the first 100 optimiser steps of C1-C3" -- smaller batches so that the fixtures stay small.  C3's
    # latent pipeline (VAE + FFJORD map) is out of scope, so its GP flow (d = 10, nb_div_blocks = 9, [50, 50, 50],
    # default_params/latent_params.py:195-198) is trained on a synthetic map output: x_map ~ N(0, I), x = a fixed affine
    # image of it plus noise; everything from `loss_gp_flow` on is the reference's own code.
    dict(name="train100_c1_eight_gaussians", data="eight_gaussians", K=100, B=256, euler=None),
    dict(name="train100_c2_moons_euler", data="moons", K=100, B=256, euler=5e-4),
    dict(name="train100_c3_latent_synthetic", data=None, K=100, B=64, euler=None, d=10, nb=9, hidden=[50, 50, 50]),
]
ONLY = [a for a in sys.argv[1:] if not a.startswith("-")]


def main():
    sys.path.insert(0, HERE)
    import torch
    import ref_import

    rn = ref_import.import_reference(fp64=False)
    from src.case import Case
    from src.gp_flows.network import Network
    from src.datasets.distributions_toy import inf_train_gen
    import gpflow_oracle as O

    for c in CASES:
        if ONLY and c["name"] not in ONLY:
            continue
        K, B, euler = c["K"], c["B"], c["euler"]
        d, nb, hidden = c.get("d", 2), c.get("nb", 1), c.get("hidden", [15, 15])
        rng = np.random.RandomState(42)
        if c["data"] is not None:
            data = inf_train_gen(c["data"], rng=rng, batch_size=K * B).astype(np.float32)
            x_all = torch.from_numpy(data).reshape(K, B, 2)
            nf = torch.load(os.path.join(REF, "pretrained_models", c["data"], "bnaf", "torch_nf_model.pt"),
                            weights_only=False, map_location="cpu")
            nf.eval()
        else:
            gen = torch.Generator().manual_seed(42)
            xm_all = torch.randn(K, B, d, generator=gen)
            A = torch.eye(d) + 0.3 * torch.randn(d, d, generator=gen) / d ** 0.5
            x_all = xm_all @ A.T + 0.1 * torch.randn(K, B, d, generator=gen) + 0.2
            nf = None
        ed = {"use_euler": euler is not None, "case": Case.penalization, "coef_penalization": euler,
              "nb_decay_penalization": 5}
        torch.manual_seed(0)
        flow = rn.get_gp_flow(d, {"nb_neurons": hidden, "nb_div_blocks": nb}, 15, euler_dict=ed)
        p0 = torch.cat([p.detach().reshape(-1) for p in flow.parameters()]).clone()
        opt = torch.optim.Adam(flow.parameters(), lr=2e-3, betas=(0.9, 0.999), weight_decay=1e-4)
        ns = types.SimpleNamespace(flow=flow, eps=O.eps_of(torch.float32), sigma=1, map=nf)
        ns.gp_flow = lambda v: Network.gp_flow(ns, v)
        xmaps, zws, ots, regs = [], [], [], []
        for k in range(K):
            x = x_all[k]
            with torch.no_grad():
                x_map = nf(x)[0] if nf is not None else xm_all[k]
            xmaps.append(x_map.clone())
            seed_zw = 5000 + k
            torch.manual_seed(seed_zw)
            zws.append(torch.stack([torch.stack([torch.randn(1, d)[0], torch.randn(1, d)[0]]) for _ in range(15)]))
            torch.manual_seed(seed_zw)                      # the reference draws the same z, w inside compute_euler_pen
            ns.map = lambda xx, _m=x_map: (_m,)
            loss_gp = Network.loss_gp_flow(ns, x)           # network.py:70-84
            reg = flow.regularization_coef()
            loss = loss_gp + reg
            opt.zero_grad()
            loss.backward()
            opt.step()
            ots.append(float(loss_gp))
            regs.append(float(reg))
            if k < 12 or k % 10 == 9:
                print("%s step %2d  ot=%.9g  reg=%.6g" % (c["name"], k, ots[-1], regs[-1]))
        pK = torch.cat([p.detach().reshape(-1) for p in flow.parameters()])
        path = os.path.join(OUT, c["name"] + "_f32.npz")
        np.savez_compressed(path, x=x_all.numpy(), x_map=torch.stack(xmaps).numpy(), zw=torch.stack(zws).numpy(),
                            params0=p0.numpy(), paramsK=pK.numpy(), ot=np.array(ots), reg=np.array(regs),
                            euler_coef=np.float64(-1.0 if euler is None else euler), K=K, B=B, d=d, nb=nb,
                            hidden=np.array(hidden))
        print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
