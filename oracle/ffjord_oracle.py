"""TEST INFRASTRUCTURE ONLY -- CPU restatement (torch, fp32/fp64) of the frozen FFJORD map either side of the GP flow.

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s CPU legs may import this file; the product path
(`gpflow_b200/ffjord.py` -> `gpflow_ffjord_integrate`) never does.

What it restates (reference file:line):
  * `ConcatSquashLinear.forward`   flow_models/ffjord/layers/diffeq_layers/basic.py:76-86
        out = (W x + b) * sigmoid(gw * t + gb) + hb * t
  * `ODEnet.forward`               flow_models/ffjord/layers/odefunc.py:160-174   (tanh between layers, none after the last;
                                   `FFJORD` builds it with layer_type="concatsquash", ffjord.py:21-28)
  * `divergence_bf`                flow_models/ffjord/layers/odefunc.py:13-17     (exact trace of d f / d y; eval mode,
                                   `FFJORD.eval_mode`, ffjord.py:58-61), here by forward-mode tangents
  * `ODEfunc.forward`              flow_models/ffjord/layers/odefunc.py:268-301   state (y, logp): d/dt = (f, -div f)
  * `CNF.forward`                  flow_models/ffjord/layers/cnf.py:44-98         times [0, T] (flipped when reverse)
  * `SequentialFlow.forward`       flow_models/ffjord/layers/container.py:14-26   blocks in order (reversed when reverse)
  * `FFJORD.forward / .backward`   flow_models/ffjord/ffjord.py:50-56             forward -> (z, -logp) with logp(0) = 0,
                                   i.e. (z, log|det dz/dx|); backward -> x only

PARITY STATUS: the vector field, its divergence, the block chaining and the sign conventions are PINNED to the unmodified
reference modules evaluated on pretrained_models/dsprites/ffjord (tests/golden/ffjord_dsprites_*.npz, made by
oracle/make_golden_ffjord.py).  The time integrator is NOT the reference's: the reference calls torchdiffeq 0.2.3
`odeint_adjoint(method="dopri5", atol=rtol=1e-5)` (cnf.py:4,66-83, requirements.txt:17), which is not vendored and not
installed here -- "parity unpinned" at that boundary.  This file (and the CUDA kernel) integrate with the classical
fixed-step RK4 (the GP flow's own scheme, src/gp_flows/ode_order_method.py:66-71); the golden vectors were produced by the
reference's CNF modules driven by that same fixed-step scheme injected in place of the missing `odeint`, and a
step-halving test bounds the integrator's own error below dopri5's 1e-5 tolerance.

Flat parameter layout (the C ABI's): for each CNF block, for each layer l = 0..L:
    [ _layer.weight (out x in, row-major) | _layer.bias (out) | _hyper_bias.weight (out) | _hyper_gate.weight (out) |
      _hyper_gate.bias (out) ]
"""
from dataclasses import dataclass, field
from typing import List, Sequence

import torch


@dataclass
class FfjordSpec:
    d: int
    hidden: Sequence[int]
    n_blocks: int = 1
    n_steps: int = 16                  # fixed RK4 steps per CNF block
    t_end: List[float] = field(default_factory=list)   # sqrt_end_time**2 per block (cnf.py:49-52); default 1.0

    def dims(self):
        return [self.d] + list(self.hidden) + [self.d]

    def block_params(self):
        n = self.dims()
        return sum(n[l] * n[l + 1] + 4 * n[l + 1] for l in range(len(n) - 1))

    def param_count(self):
        return self.n_blocks * self.block_params()

    def T(self, b):
        return float(self.t_end[b]) if self.t_end else 1.0


def flatten_state_dict(sd, prefix="model.chain."):
    """state_dict of the reference FFJORD module -> (flat, d, hidden, n_blocks, t_end)."""
    blocks = sorted({int(k[len(prefix):].split(".")[0]) for k in sd if k.startswith(prefix)})
    parts, t_end, hidden, d = [], [], None, None
    for b in blocks:
        p = "%s%d.odefunc.diffeq.layers." % (prefix, b)
        ls = sorted({int(k[len(p):].split(".")[0]) for k in sd if k.startswith(p)})
        for l in ls:
            q = "%s%d." % (p, l)
            parts += [sd[q + "_layer.weight"].reshape(-1), sd[q + "_layer.bias"].reshape(-1),
                      sd[q + "_hyper_bias.weight"].reshape(-1), sd[q + "_hyper_gate.weight"].reshape(-1),
                      sd[q + "_hyper_gate.bias"].reshape(-1)]
        widths = [int(sd["%s%d._layer.weight" % (p, l)].shape[0]) for l in ls]
        d = int(sd[p + "0._layer.weight"].shape[1])
        hidden = widths[:-1]
        t_end.append(float(sd["%s%d.sqrt_end_time" % (prefix, b)]) ** 2)
    return torch.cat([x.detach() for x in parts]), d, hidden, len(blocks), t_end


def unpack(flat, spec: FfjordSpec):
    n = spec.dims()
    out, o = [], 0
    for _ in range(spec.n_blocks):
        layers = []
        for l in range(len(n) - 1):
            i, k = n[l], n[l + 1]
            W = flat[o:o + k * i].reshape(k, i); o += k * i
            b = flat[o:o + k]; o += k
            hb = flat[o:o + k]; o += k
            gw = flat[o:o + k]; o += k
            gb = flat[o:o + k]; o += k
            layers.append((W, b, hb, gw, gb))
        out.append(layers)
    assert o == flat.numel()
    return out


def odenet(t, y, layers, want_div=False):
    """f(t, y) [B,d] and, optionally, trace(df/dy) [B] (basic.py:76-86, odefunc.py:160-174, :13-17)."""
    h = y
    B, d = y.shape
    Tn = torch.eye(d, dtype=y.dtype).expand(B, d, d) if want_div else None     # Tn[b, j, :] = d h / d y_j
    tt = torch.as_tensor(t, dtype=y.dtype)
    for l, (W, b, hb, gw, gb) in enumerate(layers):
        gate = torch.sigmoid(gw * tt + gb)
        a = (h @ W.t() + b) * gate + hb * tt
        if want_div:
            Tn = (Tn @ W.t()) * gate
        if l < len(layers) - 1:
            h = torch.tanh(a)
            if want_div:
                Tn = Tn * (1 - h * h).unsqueeze(1)
        else:
            h = a
    if want_div:
        return h, torch.diagonal(Tn, dim1=1, dim2=2).sum(-1)
    return h, None


def integrate(flat, x, spec: FfjordSpec, reverse=False, want_logdet=False):
    """FFJORD.forward (reverse=False: returns (z, logdet) or (z, None)) / FFJORD.backward (reverse=True)."""
    blocks = unpack(flat.to(x.dtype), spec)
    order = range(spec.n_blocks - 1, -1, -1) if reverse else range(spec.n_blocks)
    y = x.clone()
    logp = torch.zeros(x.shape[0], dtype=x.dtype)          # FFJORD.forward's `zero` (ffjord.py:51)
    for bi in order:
        layers = blocks[bi]
        T = spec.T(bi)
        t0, t1 = (T, 0.0) if reverse else (0.0, T)
        h = (t1 - t0) / spec.n_steps
        for k in range(spec.n_steps):
            t = t0 + k * h
            k1, d1 = odenet(t, y, layers, want_logdet)
            k2, d2 = odenet(t + 0.5 * h, y + (0.5 * h) * k1, layers, want_logdet)
            k3, d3 = odenet(t + 0.5 * h, y + (0.5 * h) * k2, layers, want_logdet)
            k4, d4 = odenet(t + h, y + h * k3, layers, want_logdet)
            y = y + (h / 6.0) * (k1 + 2 * k2 + 2 * k3 + k4)
            if want_logdet:
                logp = logp - (h / 6.0) * (d1 + 2 * d2 + 2 * d3 + d4)      # ODEfunc returns -divergence (odefunc.py:301)
    return y, (-logp if want_logdet else None)                              # ffjord.py:53: -model_x[1]
