"""TEST INFRASTRUCTURE ONLY -- golden vectors for `nb_blocks != 1` (src/gp_flows/compose.py:5-69) by EXECUTING THE
UNMODIFIED REFERENCE:  python oracle/make_golden_compose.py   (needs /root/reference; writes tests/golden/compose_*.npz)

`get_gp_flow(..., nb_blocks=2)` (res_net.py:14-67) returns a Compose of two ResNets; `Network.gp_flow` (network.py:156-189)
applies erf once, the blocks in cube space, clip / erfinv once.  Stored: the parameters of both blocks, the transported
points, log-det, OT cost, the gradients of every block, the input gradient, the reverse transport, the summed Lagrangian."""
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")
CASES = [dict(name="compose_d2_h15_b2", d=2, nb=1, hidden=[15, 15], N=5, B=16, blocks=2, std=0.3, seed=21),
         dict(name="compose_d3_nb2_b3", d=3, nb=2, hidden=[8, 6], N=4, B=12, blocks=3, std=0.25, seed=22)]


def main():
    sys.path.insert(0, HERE)
    import torch
    import ref_import
    rn = ref_import.import_reference(fp64=False)
    from src.gp_flows.network import Network
    import gpflow_oracle as O
    for c in CASES:
        d, nb, H, N, B = c["d"], c["nb"], c["hidden"], c["N"], c["B"]
        spec = O.FlowSpec(d=d, nb=nb, hidden=H, n_steps=N)
        flow = rn.get_gp_flow(d, {"nb_neurons": list(H), "nb_div_blocks": nb}, N, nb_blocks=c["blocks"])
        assert type(flow).__name__ == "Compose" and len(flow.layers) == c["blocks"]
        flats = []
        with torch.no_grad():
            for bi, layer in enumerate(flow.layers):
                ps = O.init_params(spec, c["std"], c["seed"] + 10 * bi, torch.float32)
                for p, q in zip(layer.parameters(), ps):
                    p.copy_(q)
                flats.append(O.flatten_params(ps).numpy())
        g = torch.Generator().manual_seed(100 + c["seed"])
        x_map = torch.randn(B, d, generator=g, dtype=torch.float64).float()
        x_data = (1.2 * x_map.double().flip(-1) + 0.3).float()
        xm = x_map.clone().requires_grad_(True)
        ns = types.SimpleNamespace(flow=flow, eps=O.eps_of(torch.float32), sigma=1)
        y, logdet = Network.gp_flow(ns, xm)
        lagr = flow.lagrangian.detach().clone()
        ot = ((x_data - y) ** 2).sum(-1).mean()
        (ot + flow.regularization_coef()).backward()
        grads = [torch.cat([p.grad.reshape(-1) for p in layer.parameters()]).numpy() for layer in flow.layers]
        with torch.no_grad():
            y_rev = Network.gp_flow(ns, x_data, reverse=True)
        path = os.path.join(OUT, c["name"] + "_f32.npz")
        np.savez_compressed(path, d=d, nb=nb, hidden=np.array(H), N=N, B=B, blocks=c["blocks"], x_map=x_map.numpy(),
                            x_data=x_data.numpy(), params=np.stack(flats), y=y.detach().numpy(), logdet=logdet.detach().numpy(),
                            ot=ot.detach().numpy(), lagrangian=lagr.numpy(), grads=np.stack(grads), grad_x=xm.grad.numpy(),
                            y_rev=y_rev.numpy())
        print(c["name"], "ot=%.8g lagr=%.6g |g0|=%.5g (%d bytes)" % (float(ot), float(lagr), float(np.linalg.norm(grads[0])),
                                                                   os.path.getsize(path)))


if __name__ == "__main__":
    main()
