"""Frozen FFJORD map either side of the GP flow, and the Gaussian-noise data set built from it -- SURVEY.md 8(f) rank 3.

Mirror of `flow_models/ffjord/ffjord.py:9-72` (`FFJORD(dim, n_neurons, nblocks, non_linearity="tanh")`, eval mode):

  * `forward(x)  -> (z, log_det)`   = `x_map, log_det = self.map(x)` (src/gp_flows/network.py:76-78, ffjord.py:50-53)
  * `backward(z) -> x`              = `map.backward`, the inverse used to build the Gaussian-noise training set
                                      (ffjord.py:55-56, src/data_module.py:161-186)

both as ONE CUDA launch (`gpflow_ffjord_integrate`: the whole chain of CNF blocks, fixed-step classical RK4, exact trace
for the log-determinant).  The reference integrates with torchdiffeq's adaptive dopri5 (atol = rtol = 1e-5,
layers/cnf.py:66-83); `n_steps` RK4 steps per block replace it -- 64 steps keep the pretrained dsprites map within 2e-5
of the converged solution (tests/test_ffjord_map.py), below dopri5's own tolerance.

`gaussian_noise_dataset` mirrors the Gaussian branch of `DataModule.get_dataset` (src/data_module.py:161-186) on the
device: sample the latent Gaussian, push it through `map.backward`, keep the result in HBM (the reference builds a CPU
`Dataset` through 2000-sample chunks and re-uploads every batch).

Non-tanh nonlinearities / non-concatsquash layers raise NotImplementedError (the caller keeps the reference module).
No CPU path.
"""
from __future__ import annotations

import ctypes
from typing import Optional, Sequence

import torch

from . import _lib
from .ops import _check_cuda, _ptr, _stream


class FFJORDMap(torch.nn.Module):
    def __init__(self, dim: int, n_neurons: Sequence[int], nblocks: int, flat_params: torch.Tensor,
                 t_end: Optional[Sequence[float]] = None, n_steps: int = 64):
        super().__init__()
        self.dim, self.n_neurons, self.nblocks, self.n_steps = int(dim), tuple(int(h) for h in n_neurons), int(nblocks), int(n_steps)
        if len(self.n_neurons) > _lib.FFJORD_MAX_HIDDEN or not 1 <= self.nblocks <= _lib.FFJORD_MAX_BLOCKS:
            raise NotImplementedError("gpflow_b200: FFJORD with %d hidden layers / %d blocks is not implemented"
                                      % (len(self.n_neurons), self.nblocks))
        self.t_end = tuple(float(t) for t in (t_end if t_end is not None else [1.0] * self.nblocks))
        self.register_buffer("flat", flat_params.detach().contiguous().clone())
        n = ctypes.c_int64(0)
        _lib.check(_lib.load().gpflow_ffjord_param_count(ctypes.byref(self._desc(torch.float32, False)), ctypes.byref(n)))
        if int(n.value) != self.flat.numel():
            raise ValueError("FFJORDMap: expected %d parameters, got %d" % (n.value, self.flat.numel()))

    def _desc(self, dtype, reverse, n_steps=None):
        d = _lib.GpFlowFfjordDesc(abi_version=_lib.ABI_VERSION, dtype=_lib.F64 if dtype == torch.float64 else _lib.F32,
                                  dim=self.dim, n_hidden=len(self.n_neurons), n_blocks=self.nblocks,
                                  n_steps=int(n_steps or self.n_steps), reverse=int(bool(reverse)))
        for i, h in enumerate(self.n_neurons):
            d.hidden[i] = h
        for i, t in enumerate(self.t_end):
            d.t_end[i] = t
        return d

    @staticmethod
    def from_state_dict(sd, prefix="model.chain.", n_steps: int = 64):
        """state_dict of the reference FFJORD (keys `model.chain.<b>.odefunc.diffeq.layers.<l>._layer.weight | _layer.bias |
        _hyper_bias.weight | _hyper_gate.weight | _hyper_gate.bias`, `model.chain.<b>.sqrt_end_time`)."""
        blocks = sorted({int(k[len(prefix):].split(".")[0]) for k in sd if k.startswith(prefix)})
        parts, t_end, hidden, dim = [], [], None, None
        for b in blocks:
            p = "%s%d.odefunc.diffeq.layers." % (prefix, b)
            ls = sorted({int(k[len(p):].split(".")[0]) for k in sd if k.startswith(p)})
            for l in ls:
                q = "%s%d." % (p, l)
                if q + "_hyper_gate.weight" not in sd or q + "_hyper_bias.weight" not in sd:
                    raise NotImplementedError("gpflow_b200: only ConcatSquashLinear ODEnets are implemented")
                parts += [sd[q + k].reshape(-1) for k in ("_layer.weight", "_layer.bias", "_hyper_bias.weight",
                                                           "_hyper_gate.weight", "_hyper_gate.bias")]
            widths = [int(sd["%s%d._layer.weight" % (p, l)].shape[0]) for l in ls]
            this_dim, this_hidden = int(sd[p + "0._layer.weight"].shape[1]), widths[:-1]
            if hidden is not None and (this_hidden != hidden or this_dim != dim):
                raise NotImplementedError("gpflow_b200: CNF blocks of different shapes are not implemented")
            dim, hidden = this_dim, this_hidden
            t_end.append(float(sd["%s%d.sqrt_end_time" % (prefix, b)]) ** 2)
        return FFJORDMap(dim, hidden, len(blocks), torch.cat([x.detach() for x in parts]), t_end, n_steps)

    @staticmethod
    def from_reference(module, n_steps: int = 64):
        for cnf in module.model.chain:
            fns = list(cnf.odefunc.diffeq.activation_fns)
            if any(type(f).__name__ != "Tanh" for f in fns) or getattr(cnf.odefunc, "residual", False):
                raise NotImplementedError("gpflow_b200: only tanh, non-residual ODEnets are implemented")
        return FFJORDMap.from_state_dict(module.state_dict(), n_steps=n_steps)

    def _run(self, x, reverse, want_logdet, n_steps=None):
        if not isinstance(x, torch.Tensor) or x.dtype not in (torch.float32, torch.float64):
            raise TypeError("FFJORDMap: x must be a float32 / float64 tensor")
        x = _check_cuda("x", x)
        if x.dim() != 2 or x.shape[1] != self.dim:
            raise ValueError("FFJORDMap: x must be [B, %d]" % self.dim)
        flat = self.flat if self.flat.dtype == x.dtype else self.flat.to(x.dtype)
        flat = _check_cuda("params", flat)
        B = x.shape[0]
        y = torch.empty_like(x)
        logdet = torch.empty(B, dtype=x.dtype, device=x.device) if want_logdet else None
        desc = self._desc(x.dtype, reverse, n_steps)
        with torch.cuda.device(x.device):   # launch on x's device and its current stream
            _lib.check(_lib.load().gpflow_ffjord_integrate(ctypes.byref(desc), ctypes.c_int64(B), _ptr(flat), _ptr(x),
                                                           _ptr(y), _ptr(logdet), _stream()))
        return y, logdet

    def forward(self, x, n_steps=None):
        """(z, log_det) like `FFJORD.forward` (ffjord.py:50-53)."""
        return self._run(x, False, True, n_steps)

    def map_only(self, x, n_steps=None):
        """z alone (`self.map(x)[0]`): skips the trace, ~9x cheaper at the pretrained dsprites shape."""
        return self._run(x, False, False, n_steps)[0]

    def backward(self, z, n_steps=None):
        """x like `FFJORD.backward` (ffjord.py:55-56)."""
        return self._run(z, True, False, n_steps)[0]


def gaussian_noise_dataset(map_: FFJORDMap, n_samples: int, train: bool = True, sigma: float = 1.0,
                           generator: Optional[torch.Generator] = None, device=None, dtype=torch.float32):
    """Gaussian branch of `DataModule.get_dataset` (src/data_module.py:161-186): `len = int(0.8 n_samples)` (train) or
    `int(0.2 n_samples)` points of N(0, sigma^2 I) pushed through `map.backward`, one launch, result resident on the
    device.  Returns `(x_data, x_gauss)`: the data set and the Gaussian points it came from -- the latter is the exact
    `x_map` of every sample up to the integrator's round-trip error, so a training step on this data can skip the
    forward map."""
    if not callable(getattr(map_, "backward", None)):
        raise RuntimeError("If train_gp_on_data is set to False the map should have a function backward defined.")
    n = int(0.8 * n_samples if train else 0.2 * n_samples)
    device = device if device is not None else map_.flat.device
    gauss = sigma * torch.randn(n, map_.dim, generator=generator, device=device, dtype=dtype)
    return map_.backward(gauss), gauss
