"""Drop-in mirror of `src/gp_flows/res_net.py`: `get_gp_flow` and `ResNet` on the CUDA kernels.

Same constructor arguments, attributes and methods as the reference module:
`__call__(x, training=False, current_epoch=0, reverse=False, save_trajectories=False, id_final_layer=-1)`,
`.lagrangian`, `.regularization_value`, `.euler_coef_penalization` (mutable: GPFlowCallback halves it,
gp_flow_callback.py:13-26), `.euler_dict`, `.nb_layers`, `.dt`, `.v` (`.v.v.denses`, `.v.bound`,
`.v.nb_div_blocks`), `regularization_coef()`, `get_trajectories()`, `set_nb_layers_eval(n)`, `state_dict()` with keys
`v.v.denses.{i}.weight|bias`.  Gradients arrive in `denses[i].weight.grad / .bias.grad` and in the input.

Options outside the CUDA path (spectral Euler, non-tanh activations, stationary / compressible velocities,
non-RK4 or non-standard solvers) raise NotImplementedError: an integration keeps the reference's own Python
classes for those.
"""
from __future__ import annotations

import numpy as np
import torch

from . import _lib, ops
from .ops import FlowConfig
from .velocity_field import CONTINUOUS_TIME, VelocityField


import weakref

_PROBE_RINGS = weakref.WeakKeyDictionary()   # ResNet -> pinned staging buffers of draw_probes


class Compose(torch.nn.Module):
    """compose.py:5-69 -- sequential composition of several flows (nb_blocks != 1); a thin Python loop."""

    def __init__(self, layers):
        super().__init__()
        self.layers = torch.nn.ModuleList(layers)
        self.euler_coef_penalization = self.layers[0].euler_coef_penalization

    def __call__(self, z, reverse=False, save_trajectories=False):
        if not reverse:
            return self.forward(z, save_trajectories)
        return self.backward(z, save_trajectories)

    def forward(self, z, save_trajectories=False, id_final_layer=-1):
        for layer in self.layers:
            z = layer(z, save_trajectories=save_trajectories, id_final_layer=id_final_layer)
        self.compute_lagrangian()
        return z

    def backward(self, z, save_trajectories=False):
        for layer in reversed(self.layers):
            z = layer(z, save_trajectories=save_trajectories, reverse=True)
        return z

    def get_trajectories(self):
        xs, vs = [], []
        for layer in self.layers:
            xs.extend(t.detach().cpu().numpy() for t in layer.x_trajectories)
            vs.extend(t.detach().cpu().numpy() for t in layer.v_trajectories)
        return np.array(xs), np.array(vs)

    def regularization_coef(self):
        if len(self.layers) > 1 and self.layers[0].euler_coef_penalization is not None:
            raise RuntimeError("Euler penalization only implemented for 1 num_blocks")
        return self.layers[0].regularization_value

    def compute_lagrangian(self):
        self.lagrangian = self.layers[0].lagrangian
        for l in self.layers[1:]:
            self.lagrangian = self.lagrangian + l.lagrangian
        return self.lagrangian


def get_gp_flow(dim, velocity_dict, nb_layers_gp, nb_blocks=1, T_final=1.0, ode_params={"method": "RK4"},
                euler_dict={"use_euler": False}, euler_spectral_velocity_dict=None, dtype=torch.float32):
    """res_net.py:14-67."""
    mk = lambda: ResNet(dim, velocity_dict, nb_layers_gp, T_final, ode_params, euler_dict,
                        euler_spectral_velocity_dict, dtype=dtype)
    if nb_blocks != 1:
        return Compose([mk() for _ in range(nb_blocks)])
    return mk()


class _FlowFunction(torch.autograd.Function):
    """y, reg, lagrangian = flow(x; params).  Backward re-integrates and runs the fused reverse sweep."""

    @staticmethod
    def forward(ctx, x, zw, opts, *params):
        cfg, wrap, reverse, training, euler_coef, want_traj, fg = opts
        flat = torch.cat([p.detach().reshape(-1) for p in params])
        out = ops.flow_forward(cfg, flat, x.detach(), wrap=wrap, reverse=reverse, training=training, lagrangian=True,
                               euler_coef=euler_coef, zw=zw, want_traj=want_traj, want_logdet=False, force_generic=fg)
        ctx.save_for_backward(x.detach(), flat, zw if zw is not None else x.new_empty(0))
        ctx.opts = opts
        ctx.shapes = [p.shape for p in params]
        sc = out["scalars"]
        traj = out["traj"] if want_traj else x.new_empty(0)
        ctx.mark_non_differentiable(traj)
        return out["y"], sc[_lib.SCALAR_REG].clone(), sc[_lib.SCALAR_LAGR].clone(), traj

    @staticmethod
    def backward(ctx, gy, greg, glagr, _gtraj):
        x, flat, zw = ctx.saved_tensors
        cfg, wrap, reverse, training, euler_coef, _, fg = ctx.opts
        # the Lagrangian is only ever logged (network.py:132); the reference never back-propagates it
        if gy is None:
            gy = torch.zeros_like(x)
        grads, gx = ops.flow_backward(cfg, flat, x, gy.contiguous(), wrap=wrap, reverse=reverse, training=training,
                                      euler_coef=euler_coef, zw=zw if zw.numel() else None,
                                      reg_cotangent=greg if (greg is not None and zw.numel()) else None,
                                      want_grad_x=ctx.needs_input_grad[0], force_generic=fg)
        outs, off = [], 0
        for shp in ctx.shapes:
            n = int(np.prod(shp))
            outs.append(grads[off:off + n].reshape(shp))
            off += n
        return (gx, None, None) + tuple(outs)


class ResNet(torch.nn.Module):
    def __init__(self, dim, velocity_dict, nb_layers, T_final=1.0, ode_params={"method": "RK4"},
                 euler_dict={"use_euler": False}, euler_spectral_velocity_dict=None, compute_divergence=False,
                 dtype=torch.float32):
        super().__init__()
        self.velocity_dict = velocity_dict
        self.nb_layers = nb_layers
        self.T_final = T_final
        self.dt = T_final / nb_layers
        self.set_nb_layers_eval(nb_layers)
        self.constraint_value = 0.0
        self.init_trajectories()
        self.euler_dict = euler_dict
        use_euler = bool(euler_dict.get("use_euler", False))
        case = str(euler_dict.get("case", "")) if use_euler else ""
        if use_euler and case == "penalization":
            self.euler_coef_penalization = euler_dict["coef_penalization"]
        else:
            self.euler_coef_penalization = None
        if use_euler and case == "spectral_method":
            raise NotImplementedError("gpflow_b200: euler_case spectral_method (PolyVelocity) is outside the CUDA path")
        if str(ode_params.get("method", "RK4")) != "RK4" or str(ode_params.get("solve_method", "standard")) != "standard":
            raise NotImplementedError("gpflow_b200: only RK4 / Case.standard integration is on the CUDA path")
        if compute_divergence:
            raise NotImplementedError("gpflow_b200: the debug divergence accumulator (res_net.py:304-317) is not on the CUDA path")
        self.v = VelocityField(dim, velocity_dict.get("nb_neurons", [10]),
                               activ_func=velocity_dict.get("activ_func", "tanh"),
                               time_state=velocity_dict.get("time_state", CONTINUOUS_TIME),
                               incompressible=velocity_dict.get("incompressible", True),
                               nb_div_blocks=velocity_dict.get("nb_div_blocks", 1), dtype=dtype)
        self.dim = dim
        self.compute_divergence = compute_divergence
        self.lagrangian = torch.tensor(0.0, dtype=dtype)
        self.regularization_value = torch.tensor(0.0, dtype=dtype)
        self.force_generic = False

    # -- reference API -------------------------------------------------------------------------
    def init_trajectories(self):
        self.x_trajectories = []
        self.v_trajectories = []

    def set_nb_layers_eval(self, nb_layers):
        self.nb_layers_eval = nb_layers
        self.dt_eval = self.T_final / self.nb_layers_eval

    def config(self, sigma=1.0) -> FlowConfig:
        return FlowConfig(d=self.dim, nb=self.v.nb_div_blocks, hidden=tuple(self.v.v.neurons[1:-1]),
                          n_steps=self.nb_layers, T_final=self.T_final, n_steps_eval=self.nb_layers_eval, sigma=sigma)

    def _euler_active(self, dtype):
        c = self.euler_coef_penalization
        return c is not None and c > 2 * ops.eps_of(dtype)

    def draw_probes(self, n_steps, like: torch.Tensor):
        """The reference draws z then w with torch.randn(1, d) once per step on the default (CPU) generator
        (res_net.py:330-333); same calls, same order -> same RNG stream."""
        d = self.dim
        rows = torch.empty(n_steps, 2, d)
        for i in range(n_steps):
            rows[i, 0] = torch.randn(1, d)[0]
            rows[i, 1] = torch.randn(1, d)[0]
        if like.device.type != "cuda":
            return rows.to(dtype=like.dtype)
        # host -> device without stalling the host: a pageable `.to(device)` blocks until the stream has drained, which puts
        # the host one step behind the GPU in a training loop (a gap before every kernel).  Two pinned staging buffers,
        # each guarded by the event of its last copy.
        ring = _PROBE_RINGS.setdefault(self, {"slots": [], "next": 0})   # outside the module: deepcopy / pickle stay clean
        key = (n_steps, d, like.dtype, like.device)
        if ring.get("key") != key:
            ring["key"], ring["next"] = key, 0
            ring["slots"] = [[torch.empty(n_steps, 2, d, dtype=like.dtype).pin_memory(), None] for _ in range(2)]
        slot = ring["slots"][ring["next"]]
        ring["next"] ^= 1
        if slot[1] is not None:
            slot[1].synchronize()            # the copy that last read this buffer (two draws ago) has completed
        slot[0].copy_(rows)
        out = slot[0].to(like.device, non_blocking=True)
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(like.device))
        slot[1] = ev
        return out

    def forward(self, x, training=False, current_epoch=0, reverse=False, save_trajectories=False, id_final_layer=-1,
                _wrap=False, _sigma=1.0):
        self.init_trajectories()
        cfg = self.config(_sigma)
        n = cfg.n_steps if training else cfg.n_eval
        euler = self.euler_coef_penalization if self._euler_active(x.dtype) else None
        zw = self.draw_probes(n, x) if euler is not None else None
        want_traj = save_trajectories or id_final_layer != -1
        params = list(self.parameters())
        cfg_opts = (cfg, _wrap, reverse, training, euler, want_traj, self.force_generic)
        y, reg, lagr, traj = _FlowFunction.apply(x, zw, cfg_opts, *params)
        self.regularization_value = reg
        self.lagrangian = lagr
        if save_trajectories:
            self._store_trajectories(traj, n, reverse, training)
        if id_final_layer != -1:
            if _wrap:
                raise ValueError("id_final_layer is a cube-space option")
            if torch.is_grad_enabled() and (x.requires_grad or any(q.requires_grad for q in params)):
                # the reference's x_full[id] carries gradients (res_net.py:227-233); the trajectory output of the CUDA op
                # does not, so a differentiable call integrates exactly id_final_layer steps instead
                raise NotImplementedError("gpflow_b200: id_final_layer under autograd -- call with torch.no_grad(), or "
                                          "use set_nb_layers_eval for a differentiable partial integration")
            return traj[id_final_layer]
        return y

    def _store_trajectories(self, traj, n, reverse, training):
        # process_ode_solution (res_net.py:248-272): states at i % max(nb_layers // 40, 1) == 0 or the last step,
        # and the velocity AT THE START POINT x_0 for the successive times
        cfg = self.config()
        dt = cfg.dt if training else cfg.dt_eval
        x0 = traj[0]
        t = torch.tensor(dt * self.nb_layers if reverse else 0.0, dtype=x0.dtype)
        dts = torch.tensor(-dt if reverse else dt, dtype=x0.dtype)
        sgn = -1.0 if reverse else 1.0
        self.x_trajectories = [x0]
        self.v_trajectories = [sgn * self.v(x0, float(t))]
        for i in range(n):
            t = t + dts
            if i % max(self.nb_layers // 40, 1) == 0 or i == self.nb_layers - 1:
                self.x_trajectories.append(traj[i + 1])
                self.v_trajectories.append(sgn * self.v(x0, float(t)))

    def get_trajectories(self):
        return (np.array([t.detach().cpu().numpy() for t in self.x_trajectories]),
                np.array([t.detach().cpu().numpy() for t in self.v_trajectories]))

    def regularization_coef(self):
        return self.regularization_value

    def flat_parameters(self):
        return self.v.flat_parameters()
