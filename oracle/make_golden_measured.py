"""TEST INFRASTRUCTURE ONLY -- golden vectors AT THE MEASURED HYPER-PARAMETERS (bench.py WORKLOADS: N = 15 RK4 steps,
weights N(0, w_std^2) from `gpflow_oracle.init_params(spec, w_std, seed 0)`) by EXECUTING THE UNMODIFIED REFERENCE:

    python oracle/make_golden_measured.py        (needs /root/reference; writes tests/golden/measured_*.npz)

  measured_c5_d64_nb63   d=64, nb=63, H=[256,256], N=15, w_std=0.02, B=2   (the reference needs GBs per sample here)
  measured_c3_d10_nb9    d=10, nb=9,  H=[50,50,50], N=15, w_std=0.05, B=64
The parameters are re-created from the seed by the tests (1.1 M floats would not be a small fixture); the gradient is
stored as its norm, a strided sub-sample and four random projections for C5, in full for C3."""
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")
CASES = [dict(name="measured_c3_d10_nb9", d=10, nb=9, hidden=[50, 50, 50], N=15, B=64, std=0.05, full=True),
         dict(name="measured_c5_d64_nb63", d=64, nb=63, hidden=[256, 256], N=15, B=2, std=0.02, full=False)]


def main(only=None):
    sys.path.insert(0, HERE)
    sys.path.insert(0, os.path.dirname(HERE))
    import torch
    import ref_import
    rn = ref_import.import_reference(fp64=False)
    from src.gp_flows.network import Network
    import gpflow_oracle as O
    from bench import synth_batch
    for c in CASES:
        if only and c["name"] != only:
            continue
        d, nb, H, N, B = c["d"], c["nb"], c["hidden"], c["N"], c["B"]
        spec = O.FlowSpec(d=d, nb=nb, hidden=H, n_steps=N)
        flow = rn.get_gp_flow(d, {"nb_neurons": list(H), "nb_div_blocks": nb}, N)
        ps = O.init_params(spec, c["std"], 0, torch.float32)
        with torch.no_grad():
            for p, q in zip(flow.parameters(), ps):
                p.copy_(q)
        x_data, x_map = synth_batch(B, d, 1234, "cpu")          # the bench's synthetic batch
        xm = x_map.clone().requires_grad_(True)
        ns = types.SimpleNamespace(flow=flow, eps=O.eps_of(torch.float32), sigma=1)
        y, logdet = Network.gp_flow(ns, xm)
        lagr = flow.lagrangian.detach().clone()
        ot = ((x_data - y) ** 2).sum(-1).mean()
        (ot + flow.regularization_coef()).backward()
        grads = torch.cat([p.grad.reshape(-1) for p in flow.parameters()])
        rec = dict(d=d, nb=nb, hidden=np.array(H), N=N, B=B, w_std=c["std"], w_seed=0, x_map=x_map.numpy(), x_data=x_data.numpy(),
                   y=y.detach().numpy(), logdet=logdet.detach().numpy(), ot=ot.detach().numpy(), lagrangian=lagr.numpy(),
                   grad_x=xm.grad.numpy(), grads_norm=np.float64(grads.double().norm()))
        if c["full"]:
            rec["grads"] = grads.numpy()
        else:
            idx = np.arange(0, grads.numel(), 997)
            rec["grad_idx"] = idx
            rec["grads_sub"] = grads.numpy()[idx]
            gen = np.random.default_rng(0)
            rec["grads_proj"] = np.array([float((grads.double() * torch.from_numpy(gen.standard_normal(grads.numel()))).sum())
                                          for _ in range(4)])
        path = os.path.join(OUT, c["name"] + "_f32.npz")
        np.savez_compressed(path, **rec)
        print(c["name"], "ot=%.8g lagr=%.6g |g|=%.6g (%d bytes)" % (float(ot), float(lagr), float(grads.norm()), os.path.getsize(path)),
              flush=True)


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else None)
