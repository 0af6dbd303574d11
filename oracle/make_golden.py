"""TEST INFRASTRUCTURE ONLY -- generate tests/golden/*.npz by EXECUTING THE UNMODIFIED REFERENCE.

Run in the build container (needs /root/reference):

    python oracle/make_golden.py            # spawns one subprocess per precision mode

For every case the reference's own code computes everything that is stored:
  * `get_gp_flow(...)`                         src/gp_flows/res_net.py:14-67
  * `Network.gp_flow`, `Network.loss_gp_flow`  src/gp_flows/network.py:70-84,156-189 (called
    unbound on a namespace carrying .flow/.eps/.sigma/.map -- Network.__init__ needs the
    Lightning/dataset stack which is outside the hot path)
  * `flow.regularization_coef()`, `loss.backward()`  network.py:104-111
The Euler probes z,w are whatever the reference draws (`torch.randn(1,d)` twice per step,
res_net.py:330-333) after `torch.manual_seed(case.seed_zw)`; they are re-drawn here with the same
seed and stored so the kernels can be fed identical probes.
"""
import os
import subprocess
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")

# name, d, nb, hidden, N, B, weight spec, euler_coef, extras
CASES = [
    dict(name="c1_d2_h15", d=2, nb=1, hidden=[15, 15], N=15, B=64, w=("randn", 0.3, 1), euler=None, grad_x=True, traj=True),
    dict(name="c2_d2_h15_euler", d=2, nb=1, hidden=[15, 15], N=15, B=64, w=("randn", 0.3, 1), euler=0.05),
    dict(name="d2_refinit", d=2, nb=1, hidden=[15, 15], N=15, B=64, w=("refinit", 1e-3, 3), euler=None),
    dict(name="d2_h10_L1_evalsteps", d=2, nb=1, hidden=[10], N=6, B=48, w=("randn", 0.4, 4), euler=None, n_eval=9),
    dict(name="d3_nb2_L1_euler", d=3, nb=2, hidden=[8], N=3, B=32, w=("randn", 0.3, 5), euler=0.05),
    dict(name="d5_nb3_L3_euler", d=5, nb=3, hidden=[6, 7, 4], N=4, B=32, w=("randn", 0.3, 6), euler=0.05, grad_x=True),
    dict(name="d4_nb4_L2", d=4, nb=4, hidden=[12, 9], N=5, B=32, w=("randn", 0.3, 8), euler=None),
    dict(name="c3_d10_nb9_h50", d=10, nb=9, hidden=[50, 50, 50], N=15, B=16, w=("randn", 0.05, 7), euler=None),
    dict(name="c3_d10_nb9_h50_euler", d=10, nb=9, hidden=[50, 50, 50], N=3, B=8, w=("randn", 0.05, 7), euler=5e-3),
    dict(name="c5_d64_nb63_h256", d=64, nb=63, hidden=[256, 256], N=2, B=4, w=("formula", 0.004, 0), euler=None),
]


def run(fp64: bool):
    sys.path.insert(0, HERE)
    import torch
    import ref_import

    rn = ref_import.import_reference(fp64=fp64)
    from src.case import Case
    from src.gp_flows.network import Network
    import gpflow_oracle as O

    dtype = torch.float64 if fp64 else torch.float32
    tag = "f64" if fp64 else "f32"
    os.makedirs(OUT, exist_ok=True)
    for c in CASES:
        d, nb, H, N, B = c["d"], c["nb"], c["hidden"], c["N"], c["B"]
        spec = O.FlowSpec(d=d, nb=nb, hidden=H, n_steps=N)
        euler = c["euler"]
        ed = {"use_euler": euler is not None, "case": Case.penalization, "coef_penalization": euler,
              "nb_decay_penalization": 5}
        flow = rn.get_gp_flow(d, {"nb_neurons": list(H), "nb_div_blocks": nb}, N, euler_dict=ed)
        kind, scale, seed = c["w"]
        if kind == "randn":
            ps = O.init_params(spec, scale, seed, dtype)
        elif kind == "formula":
            ps = O.formula_params(spec, scale, dtype)
        else:  # the reference's own init (vector_field.py:63-70), reseeded
            torch.manual_seed(seed)
            flow = rn.get_gp_flow(d, {"nb_neurons": list(H), "nb_div_blocks": nb}, N, euler_dict=ed)
            ps = [p.detach().clone().to(dtype) for p in flow.parameters()]
        with torch.no_grad():
            for p, q in zip(flow.parameters(), ps):
                p.copy_(q)
        assert all(p.dtype == dtype for p in flow.parameters())
        if c.get("n_eval"):
            flow.set_nb_layers_eval(c["n_eval"])
        g = torch.Generator().manual_seed(100 + seed)
        x_map = torch.randn(B, d, generator=g, dtype=torch.float64).to(dtype)
        x_data = (1.2 * x_map.double().flip(-1) + 0.3).to(dtype)
        xm = x_map.clone().requires_grad_(bool(c.get("grad_x")))
        ns = types.SimpleNamespace(flow=flow, eps=O.eps_of(dtype), sigma=1)
        ns.map = lambda xx: (xm,)
        ns.gp_flow = lambda v: Network.gp_flow(ns, v)
        seed_zw = 1000 + seed
        # --- the train step exactly as network.py:104-111 composes it
        torch.manual_seed(seed_zw)
        y, logdet = Network.gp_flow(ns, xm)
        lagr = flow.lagrangian.detach().clone()
        reg = flow.regularization_coef()
        ot = ((x_data - y) ** 2).sum(-1).mean()       # network.py:82
        loss = ot + reg
        flow.zero_grad()
        loss.backward()
        grads = torch.cat([p.grad.reshape(-1) for p in flow.parameters()])
        torch.manual_seed(seed_zw)
        zw = torch.stack([torch.stack([torch.randn(1, d)[0], torch.randn(1, d)[0]]) for _ in range(N)]).to(dtype)
        # --- raw flow (cube space) forward / reverse, as callers of ResNet.forward see it
        with torch.no_grad():
            g0 = torch.erf(x_map / np.sqrt(2.0))
            torch.manual_seed(seed_zw)
            gT = flow(g0)
            traj = None
            if c.get("traj"):
                traj = flow.ode_solver.solve(g0, torch.tensor(0.0).type_as(g0), torch.tensor(flow.dt_eval).type_as(g0),
                                             flow.nb_layers_eval)
            torch.manual_seed(seed_zw)
            y_rev = Network.gp_flow(ns, x_data, reverse=True)
            lagr_rev = flow.lagrangian.detach().clone()
        rec = dict(
            d=d, nb=nb, hidden=np.array(H), N=N, n_eval=c.get("n_eval") or N, B=B,
            euler_coef=np.float64(-1.0 if euler is None else euler),
            w_kind=kind, w_scale=scale,
            x_map=x_map.numpy(), x_data=x_data.numpy(), zw=zw.numpy(),
            y=y.detach().numpy(), logdet=logdet.detach().numpy(), g_T=gT.numpy(),
            ot=ot.detach().numpy(), reg=reg.detach().numpy(), lagrangian=lagr.numpy(),
            y_rev=y_rev.numpy(), lagrangian_rev=lagr_rev.numpy(),
        )
        if kind == "formula":
            idx = np.arange(0, grads.numel(), 997)
            rec["grad_idx"] = idx
            rec["grads_sub"] = grads.numpy()[idx]
            rec["grads_norm"] = np.float64(grads.double().norm())
            gen = np.random.default_rng(0)
            rec["grads_proj"] = np.array([float((grads.double() * torch.from_numpy(
                gen.standard_normal(grads.numel()))).sum()) for _ in range(4)])
        else:
            rec["params"] = O.flatten_params(ps).numpy()
            rec["grads"] = grads.numpy()
        if c.get("grad_x"):
            rec["grad_x"] = xm.grad.numpy()
        if traj is not None:
            rec["traj"] = traj.numpy()
        path = os.path.join(OUT, "%s_%s.npz" % (c["name"], tag))
        np.savez_compressed(path, **rec)
        print("%-28s %s ot=%.9g reg=%.6g lagr=%.6g |grad|=%.6g  (%d bytes)" % (
            c["name"], tag, float(ot), float(reg), float(lagr), float(grads.norm()), os.path.getsize(path)))


if __name__ == "__main__":
    if len(sys.argv) > 1:
        run(sys.argv[1] == "f64")
    else:
        for mode in ("f32", "f64"):
            subprocess.check_call([sys.executable, "-W", "ignore", os.path.abspath(__file__), mode])
