"""Fused Adam (+ L2) and StepLR for the GP-flow parameters -- SURVEY.md 8(f) rank 4.

Mirror of the optimiser the reference builds for the GP branch in `Network.configure_optimizers`
(src/gp_flows/network.py:307-320): `torch.optim.Adam(flow_params, lr=lr_gp_flow, betas=(0.9, 0.999),
weight_decay=weight_decay_gp)` followed by `StepLR(step_size=step_size_gp, gamma=0.5)`, and of the
`zero_grad / manual_backward / step` sequence of `training_step` (network.py:104-111).

The parameters of the flow are re-homed into ONE flat device buffer in the kernels' parameter order
`[W_1, b_1, ..., W_{L+1}, b_{L+1}]` (each `nn.Parameter` becomes a view of it, `state_dict()` keys are unchanged), and
their `.grad` tensors become views of the `[grads | ot, reg, lagr]` buffer the fused train step writes -- so one
optimisation step is: fused train kernel -> (all-reduce) -> ONE `gpflow_adam_step` launch, with no `torch.cat`, no
per-parameter `.grad` copies and no foreach optimiser launches.  At B=1024 (C1/C2) those eager ops cost more than the
0.3 ms train kernel itself.

Create it AFTER moving the module to its device / dtype (like any torch optimiser).  There is no CPU path.
"""
from __future__ import annotations

import ctypes

import torch

from . import _lib
from .ops import _dtype_code, _ptr, _stream


class FlatAdam:
    def __init__(self, params, lr=2e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, step_size=None, gamma=0.5):
        self.params = [p for p in params]
        if not self.params:
            raise ValueError("FlatAdam: empty parameter list")
        p0 = self.params[0]
        if not p0.is_cuda:
            raise RuntimeError("gpflow_b200: FlatAdam needs CUDA parameters (no CPU path exists in this package)")
        self.dtype, self.device = p0.dtype, p0.device
        self.n = sum(p.numel() for p in self.params)
        self.flat = torch.empty(self.n, dtype=self.dtype, device=self.device)
        self.gradbuf = torch.zeros(self.n + _lib.NUM_SCALARS, dtype=self.dtype, device=self.device)
        off = 0
        with torch.no_grad():
            for p in self.params:
                if p.dtype != self.dtype or p.device != self.device:
                    raise TypeError("FlatAdam: parameters must share one dtype and device")
                n = p.numel()
                self.flat[off:off + n].copy_(p.detach().reshape(-1))
                p.data = self.flat[off:off + n].view(p.shape)
                p.grad = self.gradbuf[off:off + n].view(p.shape)
                off += n
        self.exp_avg = torch.zeros(self.n, dtype=self.dtype, device=self.device)
        self.exp_avg_sq = torch.zeros(self.n, dtype=self.dtype, device=self.device)
        self.lr, self.betas, self.eps, self.weight_decay = float(lr), (float(betas[0]), float(betas[1])), float(eps), \
            float(weight_decay)
        self.step_size, self.gamma = step_size, float(gamma)
        self.step_count = 0
        self.epoch = 0

    # the flat views survive as long as nobody re-assigns `.data` / `.grad` (e.g. `module.double()`, `zero_grad(set_to_none=True)`)
    def intact(self) -> bool:
        off = 0
        es = self.flat.element_size()
        for p in self.params:
            if p.data_ptr() != self.flat.data_ptr() + off * es or p.grad is None or \
                    p.grad.data_ptr() != self.gradbuf.data_ptr() + off * es:
                return False
            off += p.numel()
        return True

    @property
    def grads(self):
        return self.gradbuf[:self.n]

    def zero_grad(self, set_to_none=False):
        self.gradbuf[:self.n].zero_()

    def step(self):
        """One torch.optim.Adam update of every parameter: a single kernel launch."""
        self.step_count += 1
        _lib.check(_lib.load().gpflow_adam_step(
            _dtype_code(self.dtype), ctypes.c_int64(self.n), _ptr(self.flat), _ptr(self.gradbuf), _ptr(self.exp_avg),
            _ptr(self.exp_avg_sq), self.lr, self.betas[0], self.betas[1], self.eps, self.weight_decay,
            ctypes.c_int64(self.step_count), _stream()))

    def scheduler_step(self):
        """StepLR(step_size, gamma) stepped once per epoch (network.py:113-114, 318-320)."""
        self.epoch += 1
        if self.step_size and self.epoch % int(self.step_size) == 0:
            self.lr *= self.gamma

    def state_dict(self):
        return {"step": self.step_count, "epoch": self.epoch, "lr": self.lr, "exp_avg": self.exp_avg.clone(),
                "exp_avg_sq": self.exp_avg_sq.clone()}

    def load_state_dict(self, sd):
        self.step_count, self.epoch, self.lr = int(sd["step"]), int(sd["epoch"]), float(sd["lr"])
        self.exp_avg.copy_(sd["exp_avg"])
        self.exp_avg_sq.copy_(sd["exp_avg_sq"])
