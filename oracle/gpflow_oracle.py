"""TEST INFRASTRUCTURE ONLY -- CPU oracle for the GP-flow (Gaussian-preserving flow) hot path.

A from-scratch restatement, in plain CPU PyTorch (fp32 or fp64, autograd for the
gradients), of the algorithm the reference implements in

  * src/gp_flows/vector_field.py:75-128      MLP potentials + forward-mode Jacobian
  * src/gp_flows/velocity_field.py:20-79     divergence-free velocity assembly
  * src/gp_flows/ode_order_method.py:66-71   classical RK4 step
  * src/gp_flows/ode_solver.py:70-81         time loop (t accumulated by repeated addition)
  * src/gp_flows/res_net.py:199-364          solve / constraints (Lagrangian, Euler penalty)
  * src/gp_flows/network.py:70-84,156-189    erf -> flow -> clip -> erfinv wrapper, OT cost

It is NOT the product: only `tests/`, `__graft_entry__.smoke()` and the `cpu_baseline` /
`--impl reference` legs of `bench.py` may import it, and only as the checker / the timed CPU
baseline.  The product path (gpflow_b200/) never imports anything from `oracle/`.

Parity pin: this file is checked against outputs of the unmodified reference executed in the
build container (`oracle/make_golden.py` -> `tests/golden/*.npz`); see
`tests/test_oracle_golden.py`.  The velocity is written in the closed form of SURVEY.md
Appendix A.2 (not the reference's masked-matmul formulation), the Euler penalty in the JVP
form of A.7 (not the reference's two functorch VJPs) -- so agreement with the golden vectors
is a real check of the restated math, not of copied code.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import List, Optional, Sequence

import numpy as np
import torch


def eps_of(dtype: torch.dtype) -> float:
    """src/precision.py:10-13 -- machine eps of the precision mode."""
    return float(np.finfo(np.float64 if dtype == torch.float64 else np.float32).eps)


@dataclass
class FlowSpec:
    """Hyper-parameters of one GP flow (res_net.py:70-176)."""

    d: int
    nb: int = 1
    hidden: Sequence[int] = (15, 15)
    n_steps: int = 15           # nb_layers (res_net.py:110)
    T_final: float = 1.0
    n_steps_eval: Optional[int] = None  # set_nb_layers_eval (res_net.py:174-176)

    @property
    def dt(self) -> float:
        return self.T_final / self.n_steps

    @property
    def n_eval(self) -> int:
        return self.n_steps if self.n_steps_eval is None else self.n_steps_eval

    @property
    def dt_eval(self) -> float:
        return self.T_final / self.n_eval

    def layer_sizes(self) -> List[int]:
        return [self.d + 1] + list(self.hidden) + [self.d * self.nb]

    def n_params(self) -> int:
        s = self.layer_sizes()
        return sum(s[i] * s[i + 1] + s[i + 1] for i in range(len(s) - 1))


def init_params(spec: FlowSpec, std: float, seed: int, dtype=torch.float64):
    """Weights AND biases ~ N(0, std^2) (the reference init N(0,1e-3)/zero-bias
    vector_field.py:63-70 gives a near-identity flow that hides errors)."""
    g = torch.Generator().manual_seed(seed)
    s = spec.layer_sizes()
    ps = []
    for i in range(len(s) - 1):
        ps.append((torch.randn(s[i + 1], s[i], generator=g, dtype=torch.float64) * std).to(dtype))
        ps.append((torch.randn(s[i + 1], generator=g, dtype=torch.float64) * std).to(dtype))
    return ps


def formula_params(spec: FlowSpec, scale: float, dtype=torch.float64):
    """Storage-free deterministic weights (for fixtures whose weights are too big to commit)."""
    s = spec.layer_sizes()
    ps = []
    for i in range(len(s) - 1):
        k = torch.arange(s[i + 1], dtype=torch.float64)[:, None]
        m = torch.arange(s[i], dtype=torch.float64)[None, :]
        ps.append((scale * torch.sin(1.3 * k + 0.7 * m + 0.5 * i + 0.1)).to(dtype))
        ps.append((scale * torch.cos(0.9 * k[:, 0] + 1.1 * i)).to(dtype))
    return ps


def flatten_params(params) -> torch.Tensor:
    """[W_1, b_1, ..., W_{L+1}, b_{L+1}], each W row-major [out, in] (vector_field.py:56-61)."""
    return torch.cat([p.reshape(-1) for p in params])


# --------------------------------------------------------------------------------------
# A.1  potentials + tangents  (vector_field.py:75-128; time column dropped, velocity_field.py:47-48)
# --------------------------------------------------------------------------------------
def potentials(params, x: torch.Tensor, t: torch.Tensor):
    """x [B,d], t 0-dim -> f [B,n_out], J [B,n_out,d] (space columns of the Jacobian only)."""
    B, d = x.shape
    n_layers = len(params) // 2
    W1, b1 = params[0], params[1]
    a = x @ W1[:, :d].T + t * W1[:, d] + b1
    if n_layers == 1:
        return a, W1[:, :d].unsqueeze(0).expand(B, -1, -1)
    h = torch.tanh(a)
    T = (1.0 - h * h).unsqueeze(-1) * W1[:, :d].unsqueeze(0)          # [B,H1,d]
    for l in range(1, n_layers - 1):
        W, b = params[2 * l], params[2 * l + 1]
        a = h @ W.T + b
        U = torch.einsum("km,bmj->bkj", W, T)
        h = torch.tanh(a)
        T = (1.0 - h * h).unsqueeze(-1) * U
    W, b = params[-2], params[-1]
    f = h @ W.T + b
    J = torch.einsum("km,bmj->bkj", W, T)
    return f, J


# --------------------------------------------------------------------------------------
# A.2  divergence-free, wall-tangent velocity (velocity_field.py:20-79), closed form
#      v_i = s_i * sum_{n<=min(i,nb-1)} sum_{j>=n} [ 2 (f_i^n - f_j^n) x_j + s_j (d_j f_i^n - d_j f_j^n) ]
# --------------------------------------------------------------------------------------
def velocity(params, x: torch.Tensor, t: torch.Tensor, nb: int, eps: float):
    B, d = x.shape
    f, J = potentials(params, x, t)
    f = f.reshape(B, d, nb)                 # f_i^n = f[:, i*nb+n]  (velocity_field.py:50)
    J = J.reshape(B, d, nb, d)              # J[b,i,n,j] = d_j f_i^n
    b = 1.0 - eps                           # velocity_field.py:96
    s = x * x - b
    idx = torch.arange(d)
    n_idx = torch.arange(nb)
    # mask[i,n,j] = (n <= i) and (j >= n)
    mask = ((n_idx[None, :, None] <= idx[:, None, None]) & (idx[None, None, :] >= n_idx[None, :, None])).to(x.dtype)
    fd = f.unsqueeze(-1) - f.permute(0, 2, 1).unsqueeze(1)            # [b,i,n,j] = f_i^n - f_j^n
    Jjj = torch.diagonal(J, dim1=1, dim2=3)                             # [b,n,j] = d_j f_j^n
    Jd = J - Jjj.unsqueeze(1)                                           # [b,i,n,j]
    Q = ((2.0 * fd * x[:, None, None, :] + s[:, None, None, :] * Jd) * mask).sum(dim=(2, 3))
    return s * Q


# --------------------------------------------------------------------------------------
# A.5  RK4 time loop  (ode_order_method.py:66-71, ode_solver.py:70-81, res_net.py:199-233)
# --------------------------------------------------------------------------------------
def _time0_dt(spec: FlowSpec, dtype, reverse: bool, training: bool):
    dt_solve = spec.dt if training else spec.dt_eval
    n = spec.n_steps if training else spec.n_eval
    t = torch.tensor(0.0, dtype=dtype)
    dt = torch.tensor(dt_solve, dtype=dtype)
    if reverse:
        t = torch.tensor(dt_solve * spec.n_steps, dtype=dtype)       # res_net.py:217-219 (nb_layers, not _eval)
        dt = -dt
    return t, dt, n


def rk4_trajectory(params, g0, spec: FlowSpec, reverse=False, training=False):
    """Returns z [n+1, B, d] and the list of step start times t_i (each a 0-dim tensor)."""
    eps = eps_of(g0.dtype)
    t, dt, n = _time0_dt(spec, g0.dtype, reverse, training)
    v = lambda x, tt: velocity(params, x, tt, spec.nb, eps)
    z = [g0]
    ts = []
    x = g0
    for _ in range(n):
        ts.append(t)
        k1 = v(x, t)
        k2 = v(x + (dt / 2.0) * k1, t + dt / 2.0)
        k3 = v(x + (dt / 2.0) * k2, t + dt / 2.0)
        k4 = v(x + dt * k3, t + dt)
        x = x + (dt / 6.0) * (k1 + 2.0 * k2 + 2.0 * k3 + k4)
        z.append(x)
        t = t + dt                                                      # ode_solver.py:80
    return torch.stack(z), ts


# --------------------------------------------------------------------------------------
# A.7  Euler penalty  (res_net.py:319-358, velocity_field.py:180-185), JVP form
# --------------------------------------------------------------------------------------
def euler_penalty_per_sample(params, x, t, z, w, nb: int):
    """q_b = z.(J_a w) - w.(J_a z), a(x) = (v(x + delta v(x,t), t+delta) - v(x,t))/delta."""
    eps = eps_of(x.dtype)
    delta = 2.0 * math.sqrt(eps)                                        # res_net.py:346

    def a(xi):
        v1 = velocity(params, xi, t, nb, eps)
        x2 = xi + delta * v1
        v2 = velocity(params, x2, t + delta, nb, eps)
        return (v2 - v1) / delta

    zb = z.expand_as(x)
    wb = w.expand_as(x)
    _, Jw = torch.func.jvp(a, (x,), (wb,))
    _, Jz = torch.func.jvp(a, (x,), (zb,))
    return (zb * Jw).sum(-1) - (wb * Jz).sum(-1)


# --------------------------------------------------------------------------------------
# res_net.py:178-317  flow forward with constraints
# --------------------------------------------------------------------------------------
@dataclass
class FlowResult:
    out: torch.Tensor
    traj: torch.Tensor
    lagrangian: torch.Tensor
    reg: torch.Tensor
    pen_steps: list = field(default_factory=list)


def flow_forward(params, g0, spec: FlowSpec, reverse=False, training=False,
                 euler_coef: Optional[float] = None, zw: Optional[torch.Tensor] = None,
                 with_lagrangian=True) -> FlowResult:
    """ResNet.forward (res_net.py:178-272).  `zw` [n,2,d]: the probes the reference draws with
    torch.randn(1,d) twice per step (z first, then w; res_net.py:330-333)."""
    dtype = g0.dtype
    eps = eps_of(dtype)
    zt, ts = rk4_trajectory(params, g0, spec, reverse, training)
    lagr = torch.tensor(0.0, dtype=dtype)
    reg = torch.tensor(0.0, dtype=dtype)
    pens = []
    n = zt.shape[0] - 1
    use_pen = euler_coef is not None and euler_coef > 2 * eps         # res_net.py:295-299
    for i in range(n):
        x = zt[i + 1]
        t = ts[i]                                                       # pairs x_{i+1} with t_i (res_net.py:259-262)
        if with_lagrangian:
            vv = velocity(params, x, t, spec.nb, eps)
            lagr = lagr + spec.dt * (0.5 * vv ** 2).sum(-1).mean()     # always self.dt (res_net.py:293)
        if use_pen:
            q = euler_penalty_per_sample(params, x, t, zw[i, 0:1].to(dtype), zw[i, 1:2].to(dtype), spec.nb)
            pen = (q ** 2).mean()
            pens.append(pen)
            reg = reg + euler_coef * pen
    return FlowResult(out=zt[-1], traj=zt, lagrangian=lagr, reg=reg, pen_steps=pens)


# --------------------------------------------------------------------------------------
# A.6  network.py:156-189 wrapper and network.py:70-84 OT-cost loss
# --------------------------------------------------------------------------------------
def gp_flow(params, x_map, spec: FlowSpec, reverse=False, sigma=1.0, **kw):
    eps = eps_of(x_map.dtype)
    g = torch.erf(x_map / (math.sqrt(2.0) * sigma))
    res = flow_forward(params, g, spec, reverse=reverse, **kw)
    y = math.sqrt(2.0) * sigma * torch.erfinv(torch.clip(res.out, -1 + eps, 1 - eps))
    logdet = ((y ** 2).sum(1) - (x_map ** 2).sum(1)) / (2.0 * sigma ** 2)
    return y, logdet, res


def train_step(params, x_data, x_map, spec: FlowSpec, euler_coef=None, zw=None,
               with_lagrangian=False, need_grad_x=False):
    """loss_gp_flow + regularization_coef + backward (network.py:70-114, GP branch).
    Returns a dict of detached tensors: y, logdet, ot, reg, loss, lagrangian, grads (flat)."""
    ps = [p.detach().clone().requires_grad_(True) for p in params]
    xm = x_map.detach().clone().requires_grad_(need_grad_x)
    y, logdet, res = gp_flow(ps, xm, spec, euler_coef=euler_coef, zw=zw, with_lagrangian=with_lagrangian)
    ot = ((x_data - y) ** 2).sum(-1).mean()
    loss = ot + res.reg
    loss.backward()
    out = dict(
        y=y.detach(), logdet=logdet.detach(), g_T=res.out.detach(), ot=ot.detach(), reg=res.reg.detach(),
        loss=loss.detach(), lagrangian=res.lagrangian.detach(),
        grads=torch.cat([(p.grad if p.grad is not None else torch.zeros_like(p)).reshape(-1) for p in ps]),
    )
    if need_grad_x:
        out["grad_x"] = xm.grad.detach()
    return out


def algorithmic_flops_per_sample(spec: FlowSpec, train=True) -> float:
    """pruned flop count per sample (FMA = 2 flops); train step = 3 x forward."""
    d, nb = spec.d, spec.nb
    H = list(spec.hidden)
    n = [d + 1] + H
    R = sum(d - k for k in range(nb))
    R2 = sum((d - k) ** 2 for k in range(nb))
    f_val = sum(2 * n[l] * n[l + 1] for l in range(len(n) - 1)) + 2 * H[-1] * R
    f_jac = H[0] * d + sum(2 * H[l - 1] * H[l] * d + H[l] * d for l in range(1, len(H))) + 2 * H[-1] * R2
    f_asm = 6 * R2
    f_eval = f_val + f_jac + f_asm
    fwd = 4 * spec.n_steps * f_eval
    return float(3 * fwd if train else fwd)
