"""TEST INFRASTRUCTURE ONLY -- golden numbers of `compute_ot_costs` (src/plots/outputs.py:7-248) by EXECUTING THE
UNMODIFIED REFERENCE on the pretrained C1 map:  python oracle/make_golden_ot_costs.py  (needs /root/reference)

net.map = pretrained_models/eight_gaussians/bnaf (BNAF: no `backward`, so the inverse-consistency block is skipped, as in
the reference's own C1 run); net.flow = get_gp_flow(2, [15, 15], 15) with seeded N(0, 0.3^2) parameters; x_test = 3000
points of the toy generator (`inf_train_gen`, seed 11).  The report's numbers are recovered from the text the reference
writes to ot_costs.txt and, unrounded, by re-running its formulas on the tensors it computes."""
import io
import os
import re
import sys
import types
from contextlib import redirect_stdout

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")
REF = "/root/reference"


def main():
    sys.path.insert(0, HERE)
    import torch
    import ref_import
    rn = ref_import.import_reference(fp64=False)
    from src.gp_flows.network import Network
    from src.gp_flows.probability_distribution import ProbabilityDistribution
    from src.datasets.distributions_toy import inf_train_gen
    sys.modules.setdefault("matplotlib", types.ModuleType("matplotlib"))
    try:
        from src.plots.outputs import compute_ot_costs
    except Exception:   # plots/__init__ pulls matplotlib: load the module file alone
        import importlib.util
        tb = types.ModuleType("src.plots.plot_tool_box")

        def make_meshgrid(bounds, Nx=50):
            lsp = [np.linspace(*bi, Nx, endpoint=True) for bi in bounds]
            xy = np.meshgrid(*lsp)
            return np.c_[xy[0].ravel(), xy[1].ravel()], xy[0], xy[1]
        tb.make_meshgrid = make_meshgrid
        pk = types.ModuleType("src.plots"); pk.__path__ = [os.path.join(REF, "src", "plots")]
        sys.modules["src.plots"] = pk
        sys.modules["src.plots.plot_tool_box"] = tb
        spec = importlib.util.spec_from_file_location("src.plots.outputs", os.path.join(REF, "src", "plots", "outputs.py"))
        mod = importlib.util.module_from_spec(spec); sys.modules["src.plots.outputs"] = mod
        spec.loader.exec_module(mod)
        compute_ot_costs = mod.compute_ot_costs
    import gpflow_oracle as O

    nf = torch.load(os.path.join(REF, "pretrained_models", "eight_gaussians", "bnaf", "torch_nf_model.pt"), weights_only=False,
                    map_location="cpu")
    nf.eval()
    spec = O.FlowSpec(d=2, nb=1, hidden=[15, 15], n_steps=15)
    flow = rn.get_gp_flow(2, {"nb_neurons": [15, 15]}, 15)
    ps = O.init_params(spec, 0.3, 31, torch.float32)
    with torch.no_grad():
        for p, q in zip(flow.parameters(), ps):
            p.copy_(q)
    x_test = torch.from_numpy(inf_train_gen("eight_gaussians", rng=np.random.RandomState(11), batch_size=3000).astype(np.float32))
    net = types.SimpleNamespace(flow=flow, map=nf, eps=O.eps_of(torch.float32), sigma=1, target_gaussian=True,
                                data=types.SimpleNamespace(dim=2, batch_size=1024, use_backward=True),
                                probability_distribution=ProbabilityDistribution(2))
    net.gp_flow = lambda *a, **k: Network.gp_flow(net, *a, **k)
    net.ot_map = lambda *a, **k: Network.ot_map(net, *a, **k)
    buf = io.StringIO()
    torch.manual_seed(77)
    with redirect_stdout(buf):
        compute_ot_costs(net, x_test, logger=None, output_dir=None, device=None)
    text = buf.getvalue()
    nums = {k: float(v) for k, v in re.findall(r"- ([A-Za-z_ ]+?) = ([-0-9.]+)", text)}
    # unrounded values with the reference's own modules / formulas
    with torch.no_grad():
        y_map, ld_map = nf(x_test)
        y_ot, ld_ot = Network.ot_map(net, x_test, return_log_det=True)
        pd = net.probability_distribution
        nll_map = float(-pd.log_prob(y_map).mean() - ld_map.mean())
        nll_ot = float(-pd.log_prob(y_ot).mean() - ld_ot.mean())
        cost_map = float(((y_map - x_test) ** 2).sum(-1).mean())
        cost_gp = float(((y_ot - x_test) ** 2).sum(-1).mean())
        ls = np.linspace(-1.0, 1.0, 100, endpoint=True)
        xx, yy = np.meshgrid(ls, ls)
        mesh = torch.tensor(np.c_[xx.ravel(), yy.ravel()]).float()
        lag = []
        for xi in torch.split(mesh, 1024):
            flow(xi)
            lag.append(float(flow.lagrangian))
        lagr = float(np.mean(lag))
    path = os.path.join(OUT, "ot_costs_c1_f32.npz")
    np.savez_compressed(path, x_test=x_test.numpy(), flow_params=O.flatten_params(ps).numpy(), report=np.array(text),
                        nll_map=nll_map, nll_ot=nll_ot, cost_map=cost_map, cost_gp_map=cost_gp, lagrangian=lagr,
                        y_ot=y_ot.numpy(), log_det_ot=ld_ot.numpy())
    print(text)
    print("unrounded:", nll_map, cost_map, nll_ot, cost_gp, lagr, "| parsed:", nums, "|", os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
