"""`compute_ot_costs` (src/plots/outputs.py:7-248) on the CUDA path -- the report `save_outputs` prints after training
(src/training.py:36-75): NLL of the map and of the OT map, OT costs, the GP-flow Lagrangian on the 100 x 100 mesh of
[-1, 1]^2 (or on uniform cube samples for d > 2), and the inverse-consistency errors when the map has a `backward`.

Same batching as the reference's `apply_fn_batch` (src/tool_box.py:57-75: `torch.split(x, batch_size)`, results
concatenated), same reductions, same text.  Everything heavy runs in the kernels behind `net.map`, `net.ot_map`,
`net.flow` (GPFlowNetwork / ResNet mirrors); this file is the host-side composition of those calls.
"""
import math

import numpy as np
import torch


def _apply_fn_batch(fn, x, batch_size):
    out = None
    for xi in torch.split(x, batch_size):
        o = fn(xi)
        if out is None:
            out = list(o) if isinstance(o, (list, tuple)) else o
        elif isinstance(o, (list, tuple)):
            out = [torch.cat((a, b)) for a, b in zip(out, o)]
        else:
            out = torch.cat((out, o))
    return out


def _log_prob_std_normal(x):
    """ProbabilityDistribution(dim).log_prob (src/gp_flows/probability_distribution.py:5-33): N(0, I)."""
    return -0.5 * (x ** 2).sum(-1) - 0.5 * x.shape[1] * math.log(2.0 * math.pi)


def make_meshgrid(n=100):
    """plots/plot_tool_box.py:88-113 with bounds ([-1, 1], [-1, 1]): points (n*n, 2), x fastest."""
    ls = np.linspace(-1.0, 1.0, n, endpoint=True)
    xx, yy = np.meshgrid(ls, ls)
    return np.c_[xx.ravel(), yy.ravel()]


@torch.no_grad()
def compute_ot_costs(net, x_test, batch_size, x_gauss=None, use_backward=True, only_ot_costs=False):
    """Returns (results dict of python floats, report string).  `x_gauss`: the N(0, I) samples the reference draws with
    `net.probability_distribution.sample([n])` (outputs.py:46); drawn here if not given."""
    dim = x_test.shape[1]
    dev = x_test.device
    map_only = net.flow is None
    res = {}
    y_map, log_det_map = _apply_fn_batch(lambda v: net.map(v), x_test, batch_size)
    if not map_only:
        y_ot, log_det_ot = _apply_fn_batch(lambda v: net.ot_map(v, return_log_det=True), x_test, batch_size)
    if x_gauss is None:
        x_gauss = torch.randn(x_test.shape[0], dim, device=dev, dtype=x_test.dtype)
    if not map_only:
        def flow_lagrangian(v):
            net.flow(v)
            return net.flow.lagrangian.reshape(1).to(v.dtype)

        if dim == 2:     # uniform mesh of the square (outputs.py:57-69)
            pts = torch.from_numpy(make_meshgrid(100)).to(device=dev, dtype=x_gauss.dtype)
        else:            # uniform samples of the hypercube (outputs.py:70-79)
            pts = (2 * (torch.rand(x_gauss.shape, device=dev) - 0.5)).to(x_test.dtype)
        res["lagrangian"] = float(_apply_fn_batch(flow_lagrangian, pts, batch_size).mean())
    has_backward = use_backward and callable(getattr(net.map, "backward", None))
    if has_backward:
        y_map_back = _apply_fn_batch(net.map.backward, x_gauss, batch_size)
        finv_f = _apply_fn_batch(lambda v: net.map(v), y_map_back, batch_size)[0]
        res["finv_f_err"] = float((finv_f - x_gauss).norm(dim=1).mean())
        res["cost_back"] = float(((y_map_back - x_gauss) ** 2).sum(-1).mean())
        if not map_only:
            y_gp_back = _apply_fn_batch(lambda v: net.ot_map(v, reverse=True), x_gauss, batch_size)
            finv_f_ot = _apply_fn_batch(net.ot_map, y_gp_back, batch_size)
            res["finv_f_gp_err"] = float((finv_f_ot - x_gauss).norm(dim=1).mean())
            res["cost_gp_back"] = float(((y_gp_back - x_gauss) ** 2).sum(-1).mean())
    res["nll_map"] = float(-_log_prob_std_normal(y_map).mean() - log_det_map.mean())
    res["cost_map"] = float(((y_map - x_test) ** 2).sum(-1).mean())
    if not map_only:
        res["nll_ot"] = float(-_log_prob_std_normal(y_ot).mean() - log_det_ot.mean())
        res["cost_gp_map"] = float(((y_ot - x_test) ** 2).sum(-1).mean())
    # the report text, line for line as outputs.py:180-240 writes it to ot_costs.txt
    nl = "  \n  "
    s = nl + "  ** Standard map: ** " + nl
    if not only_ot_costs:
        s += "- Loss map test dataset = " + "{:.3f}".format(res["nll_map"]) + nl
    s += "- OT cost map test dataset = " + "{:.3f}".format(res["cost_map"]) + nl
    if has_backward:
        s += "- OT cost map backward random data = " + "{:.3f}".format(res["cost_back"]) + nl
        s += "- finv_f error map = " + "{:.5f}".format(res["finv_f_err"]) + nl
    if not map_only:
        s += "  ** GP map: ** " + nl
        if not only_ot_costs:
            s += "- Loss GP test dataset = " + "{:.3f}".format(res["nll_ot"]) + nl
        s += "- OT cost GP test dataset = " + "{:.3f}".format(res["cost_gp_map"]) + nl
        if has_backward:
            s += "- OT cost GP backward random data = " + "{:.3f}".format(res["cost_gp_back"]) + nl
            s += "- finv_f error GP map = " + "{:.5f}".format(res["finv_f_gp_err"]) + nl
        s += "- GP flow lagrangian = " + "{:.3f}".format(res["lagrangian"]) + nl
    return res, s
