"""Frozen BNAF map in front of the GP flow -- SURVEY.md 8(f) rank 2.

Mirror of the forward of `flow_models/bnaf/bnaf.py:11-39` (`BNAF(dim, n_neurons, nb_layers, nblocks=1)`), the
pretrained normalising-flow map `Network.loss_gp_flow` evaluates under `no_grad` (`x_map = self.map(x)[0]`,
src/gp_flows/network.py:76-78).  `BNAFMap.from_reference(module)` packs the `MaskedWeight` parameters of a loaded
reference module (pretrained_models/<data>/bnaf/torch_nf_model.pt) into one flat buffer; `forward(x)` returns
`(y, log_abs_det)` like the reference module, computed by ONE CUDA launch (`gpflow_bnaf_forward`).

Shapes the kernel does not implement (more than one block, dim != 2, n_neurons / dim != 16) raise
NotImplementedError: the caller keeps the reference module for those.  BNAF has no `.backward` (inverse) in the
reference either.  No CPU path.
"""
from __future__ import annotations

import ctypes

import torch

from . import _lib
from .ops import _check_cuda, _ptr, _stream


class BNAFMap(torch.nn.Module):
    def __init__(self, dim: int, n_neurons: int, n_masked_layers: int, flat_params: torch.Tensor):
        super().__init__()
        if n_neurons % dim:
            raise ValueError("n_neurons must be a multiple of dim")
        self.dim, self.n_neurons, self.n_masked_layers = int(dim), int(n_neurons), int(n_masked_layers)
        self.register_buffer("flat", flat_params.detach().to(torch.float32).contiguous().clone())
        self._desc = _lib.GpFlowBnafDesc(abi_version=_lib.ABI_VERSION, dtype=_lib.F32, dim=self.dim,
                                         hidden_per_dim=self.n_neurons // self.dim, n_layers=self.n_masked_layers)
        n = ctypes.c_int64(0)
        rc = _lib.load().gpflow_bnaf_param_count(ctypes.byref(self._desc), ctypes.byref(n))
        if rc == _lib.ERR_UNSUPPORTED:
            raise NotImplementedError("gpflow_b200: BNAF shape dim=%d n_neurons=%d is not implemented by the CUDA kernel"
                                      % (dim, n_neurons))
        _lib.check(rc)
        if int(n.value) != self.flat.numel():
            raise ValueError("BNAFMap: expected %d parameters, got %d" % (n.value, self.flat.numel()))

    @staticmethod
    def from_state_dict(sd, prefix="flow.0."):
        """state_dict of the reference BNAF (keys `flow.0.<2l>._weight | _diag_weight | bias`, bnaf.py:21-33)."""
        if any(k.startswith("flow.1.") for k in sd):
            raise NotImplementedError("gpflow_b200: BNAF with more than one block (gated residual + flip) is not implemented")
        idx = sorted({int(k[len(prefix):].split(".")[0]) for k in sd if k.startswith(prefix) and k.endswith("._weight")})
        parts = []
        for i in idx:
            w = sd["%s%d._weight" % (prefix, i)]
            parts += [w.reshape(-1), sd["%s%d._diag_weight" % (prefix, i)].reshape(-1), sd["%s%d.bias" % (prefix, i)].reshape(-1)]
        first = sd["%s%d._weight" % (prefix, idx[0])]
        dim, n_neurons = int(first.shape[1]), int(first.shape[0])
        return BNAFMap(dim, n_neurons, len(idx), torch.cat([p.detach().float() for p in parts]))

    @staticmethod
    def from_reference(module):
        return BNAFMap.from_state_dict(module.state_dict())

    def forward(self, x):
        x = _check_cuda("x", x, torch.float32)
        if x.dim() != 2 or x.shape[1] != self.dim:
            raise ValueError("BNAFMap: x must be [B, %d]" % self.dim)
        B = x.shape[0]
        y = torch.empty_like(x)
        logdet = torch.empty(B, dtype=torch.float32, device=x.device)
        flat = self.flat if self.flat.device == x.device else self.flat.to(x.device)
        with torch.cuda.device(x.device):   # launch on x's device and its current stream
            _lib.check(_lib.load().gpflow_bnaf_forward(ctypes.byref(self._desc), ctypes.c_int64(B), _ptr(flat), _ptr(x),
                                                       _ptr(y), _ptr(logdet), _stream()))
        return y, logdet
