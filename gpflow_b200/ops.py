"""Tensor-level entry points: torch tensors in, C-ABI calls (raw device pointers + stream) out.

PyTorch is plumbing here (device memory via the caching allocator, the current stream); all arithmetic of
the GP-flow path runs in the CUDA library.  Every function raises on CPU tensors -- there is no fallback.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib
from ._lib import GpFlowDesc


@dataclass(frozen=True)
class FlowConfig:
    """Hyper-parameters of one GP flow, as `get_gp_flow` receives them (res_net.py:14-67)."""

    d: int
    nb: int = 1
    hidden: Tuple[int, ...] = (15, 15)
    n_steps: int = 15            # nb_layers
    T_final: float = 1.0
    n_steps_eval: Optional[int] = None   # nb_layers_eval (res_net.py:174-176)
    sigma: float = 1.0           # Network.sigma (network.py:51)

    def __post_init__(self):
        object.__setattr__(self, "hidden", tuple(int(h) for h in self.hidden))

    @property
    def dt(self) -> float:
        return self.T_final / self.n_steps

    @property
    def n_eval(self) -> int:
        return self.n_steps if self.n_steps_eval is None else self.n_steps_eval

    @property
    def dt_eval(self) -> float:
        return self.T_final / self.n_eval

    def layer_sizes(self):
        return [self.d + 1] + list(self.hidden) + [self.d * self.nb]

    @property
    def n_params(self) -> int:
        s = self.layer_sizes()
        return sum(s[i] * s[i + 1] + s[i + 1] for i in range(len(s) - 1))


def eps_of(dtype: torch.dtype) -> float:
    """src/precision.py:10-13."""
    return float(np.finfo(np.float64 if dtype == torch.float64 else np.float32).eps)


def _dtype_code(dtype):
    if dtype == torch.float32:
        return _lib.F32
    if dtype == torch.float64:
        return _lib.F64
    raise TypeError("gpflow_b200 supports float32 and float64 (the reference's two precision modes), got %s" % dtype)


def make_desc(cfg: FlowConfig, dtype, *, wrap=False, reverse=False, training=False, lagrangian=False,
              euler_coef: Optional[float] = None, loss_scale: float = 1.0, force_generic=False,
              force_v1=False, force_tc=False) -> GpFlowDesc:
    d = GpFlowDesc()
    d.abi_version = _lib.GPFLOW_ABI_VERSION
    d.dtype = _dtype_code(dtype)
    d.d, d.nb, d.n_hidden = cfg.d, cfg.nb, len(cfg.hidden)
    if len(cfg.hidden) > _lib.GPFLOW_MAX_HIDDEN:
        raise ValueError("at most %d hidden layers" % _lib.GPFLOW_MAX_HIDDEN)
    for i, h in enumerate(cfg.hidden):
        d.hidden[i] = h
    # ResNet.solve: training -> (dt, nb_layers) else (dt_eval, nb_layers_eval)  (res_net.py:209-214)
    d.n_steps = cfg.n_steps if training else cfg.n_eval
    d.n_steps_nominal = cfg.n_steps
    d.dt = cfg.dt if training else cfg.dt_eval
    d.dt_constraint = cfg.dt
    d.sigma = cfg.sigma
    use_pen = euler_coef is not None and euler_coef > 2 * eps_of(dtype)      # res_net.py:295-299
    d.euler_coef = float(euler_coef) if use_pen else 0.0
    d.loss_scale = float(loss_scale)
    flags = 0
    if wrap:
        flags |= _lib.FLAG_WRAP_GAUSS
    if lagrangian:
        flags |= _lib.FLAG_LAGRANGIAN
    if use_pen:
        flags |= _lib.FLAG_EULER_PEN
    if reverse:
        flags |= _lib.FLAG_REVERSE
    if force_generic:
        flags |= _lib.FLAG_FORCE_GENERIC
    if force_v1:
        flags |= _lib.FLAG_FORCE_V1
    if force_tc:
        flags |= _lib.FLAG_FORCE_TC
    d.flags = flags
    return d


# kernels enqueued through CUDA-graph replays of the tensor-core path (GPFlowNetwork._graphed_train_step): the library's own
# counter (gpflow_tc_launches_total) only sees host-side launches, i.e. the capture
REPLAYED_TC_LAUNCHES = [0]


def tc_launches_total() -> int:
    """Kernels the tensor-core path has enqueued in this process: eager launches + launches inside replayed graphs."""
    return int(_lib.load().gpflow_tc_launches_total()) + REPLAYED_TC_LAUNCHES[0]


def kernel_variant(cfg: FlowConfig, dtype=torch.float32, op=_lib.OP_TRAIN_STEP, **kw) -> str:
    desc = make_desc(cfg, dtype, **kw)
    buf = ctypes.create_string_buffer(256)
    _lib.check(_lib.load().gpflow_kernel_variant(ctypes.byref(desc), op, buf, 256))
    return buf.value.decode()


def launch_count(cfg: FlowConfig, dtype=torch.float32, op=_lib.OP_TRAIN_STEP) -> int:
    desc = make_desc(cfg, dtype)
    return int(_lib.load().gpflow_launch_count(ctypes.byref(desc), op))


def _ptr(t: Optional[torch.Tensor]):
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


def _check_cuda(name, t: torch.Tensor, dtype=None, shape=None):
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise RuntimeError("gpflow_b200: `%s` must be a CUDA tensor (no CPU path exists in this package)" % name)
    if dtype is not None and t.dtype != dtype:
        raise TypeError("gpflow_b200: `%s` has dtype %s, expected %s" % (name, t.dtype, dtype))
    if shape is not None and tuple(t.shape) != tuple(shape):
        raise ValueError("gpflow_b200: `%s` has shape %s, expected %s" % (name, tuple(t.shape), tuple(shape)))
    return t.contiguous()


def _workspace(desc, B, op, device):
    n = ctypes.c_size_t(0)
    _lib.check(_lib.load().gpflow_query_workspace(ctypes.byref(desc), B, op, ctypes.byref(n)))
    ws = torch.empty(max(int(n.value), 16), dtype=torch.uint8, device=device)
    return ws, int(n.value)


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _prep_zw(zw, desc, cfg, dtype, device):
    if not (desc.flags & _lib.FLAG_EULER_PEN):
        return None
    if zw is None:
        raise ValueError("Euler penalization needs the probes zw [n_steps, 2, d] (res_net.py:330-333)")
    zw = zw.to(device=device, dtype=dtype)
    return _check_cuda("zw", zw, dtype, (desc.n_steps, 2, cfg.d))


def flow_forward(cfg: FlowConfig, params: torch.Tensor, x: torch.Tensor, *, wrap=False, reverse=False,
                 training=False, lagrangian=False, euler_coef=None, zw=None, want_logdet=False, want_traj=False,
                 want_scalars=True, loss_scale: Optional[float] = None, force_generic=False, force_v1=False, force_tc=False):
    """ResNet.forward (raw, cube space) or Network.gp_flow (wrap=True).  Returns dict(y, logdet, traj, scalars)."""
    x = _check_cuda("x", x)
    dtype = x.dtype
    B = x.shape[0]
    x = _check_cuda("x", x, dtype, (B, cfg.d))
    params = _check_cuda("params", params, dtype, (cfg.n_params,))
    desc = make_desc(cfg, dtype, wrap=wrap, reverse=reverse, training=training, lagrangian=lagrangian,
                     euler_coef=euler_coef, loss_scale=(1.0 / max(B, 1)) if loss_scale is None else loss_scale,
                     force_generic=force_generic, force_v1=force_v1, force_tc=force_tc)
    zw = _prep_zw(zw, desc, cfg, dtype, x.device)
    y = torch.empty_like(x)
    logdet = torch.empty(B, dtype=dtype, device=x.device) if (want_logdet and wrap) else None
    traj = torch.empty(desc.n_steps + 1, B, cfg.d, dtype=dtype, device=x.device) if want_traj else None
    scalars = torch.zeros(_lib.NUM_SCALARS, dtype=dtype, device=x.device) if want_scalars else None
    with torch.cuda.device(x.device):
        ws, n = _workspace(desc, B, _lib.OP_FORWARD, x.device)
        if B > 0:
            _lib.check(_lib.load().gpflow_forward(ctypes.byref(desc), B, _ptr(params), _ptr(x), _ptr(zw), _ptr(y),
                                                  _ptr(logdet), _ptr(traj), _ptr(scalars), _ptr(ws), n, _stream()))
    return dict(y=y, logdet=logdet, traj=traj, scalars=scalars)


def fused_train_step(cfg: FlowConfig, params: torch.Tensor, x_data: torch.Tensor, x_map: torch.Tensor, *,
                     euler_coef=None, zw=None, lagrangian=False, loss_scale: Optional[float] = None, want_y=False,
                     want_grad_x=False, out: Optional[torch.Tensor] = None, force_generic=False, force_v1=False, force_tc=False):
    """loss_gp_flow + regularization_coef + backward of the GP branch (network.py:70-84,104-111) in one fused call.

    Returns dict(buf, grads, scalars, y, grad_x).  `buf` is ONE contiguous tensor [P + 3] = [grads | ot, reg, lagr]
    (views `grads`, `scalars`), already scaled by loss_scale (default 1/B) so that a single all-reduce(SUM) over
    data-parallel ranks -- with loss_scale = 1/B_global -- yields the global-batch loss and gradient.
    """
    x_map = _check_cuda("x_map", x_map)
    dtype = x_map.dtype
    B = x_map.shape[0]
    x_map = _check_cuda("x_map", x_map, dtype, (B, cfg.d))
    x_data = _check_cuda("x_data", x_data, dtype, (B, cfg.d))
    params = _check_cuda("params", params, dtype, (cfg.n_params,))
    desc = make_desc(cfg, dtype, wrap=True, training=False, lagrangian=lagrangian, euler_coef=euler_coef,
                     loss_scale=(1.0 / max(B, 1)) if loss_scale is None else loss_scale, force_generic=force_generic,
                     force_v1=force_v1, force_tc=force_tc)
    zw = _prep_zw(zw, desc, cfg, dtype, x_map.device)
    P = cfg.n_params
    if out is None:
        out = torch.empty(P + _lib.NUM_SCALARS, dtype=dtype, device=x_map.device)
    else:
        out = _check_cuda("out", out, dtype, (P + _lib.NUM_SCALARS,))
    grads, scalars = out[:P], out[P:]
    y = torch.empty_like(x_map) if want_y else None
    gx = torch.empty_like(x_map) if want_grad_x else None
    with torch.cuda.device(x_map.device):
        ws, n = _workspace(desc, B, _lib.OP_TRAIN_STEP, x_map.device)
        if B > 0:
            _lib.check(_lib.load().gpflow_train_step(ctypes.byref(desc), B, _ptr(params), _ptr(x_data), _ptr(x_map),
                                                     _ptr(zw), _ptr(grads), _ptr(scalars), _ptr(y), _ptr(gx), _ptr(ws),
                                                     n, _stream()))
        else:
            out.zero_()
    return dict(buf=out, grads=grads, scalars=scalars, y=y, grad_x=gx)


def flow_backward(cfg: FlowConfig, params: torch.Tensor, x_in: torch.Tensor, grad_y: torch.Tensor, *, wrap=False,
                  reverse=False, training=False, euler_coef=None, zw=None, reg_cotangent: Optional[torch.Tensor] = None,
                  loss_scale: Optional[float] = None, want_grad_x=True, force_generic=False, force_v1=False, force_tc=False):
    """Reverse sweep for an arbitrary loss on top of the flow: returns (grads [P], grad_x [B,d] or None)."""
    x_in = _check_cuda("x_in", x_in)
    dtype = x_in.dtype
    B = x_in.shape[0]
    x_in = _check_cuda("x_in", x_in, dtype, (B, cfg.d))
    grad_y = _check_cuda("grad_y", grad_y, dtype, (B, cfg.d))
    params = _check_cuda("params", params, dtype, (cfg.n_params,))
    desc = make_desc(cfg, dtype, wrap=wrap, reverse=reverse, training=training, euler_coef=euler_coef,
                     loss_scale=(1.0 / max(B, 1)) if loss_scale is None else loss_scale, force_generic=force_generic,
                     force_v1=force_v1, force_tc=force_tc)
    zw = _prep_zw(zw, desc, cfg, dtype, x_in.device)
    if reg_cotangent is not None:
        reg_cotangent = _check_cuda("reg_cotangent", reg_cotangent.reshape(1).to(dtype), dtype, (1,))
    grads = torch.empty(cfg.n_params, dtype=dtype, device=x_in.device)
    gx = torch.empty_like(x_in) if want_grad_x else None
    with torch.cuda.device(x_in.device):
        ws, n = _workspace(desc, B, _lib.OP_BACKWARD, x_in.device)
        if B > 0:
            _lib.check(_lib.load().gpflow_backward(ctypes.byref(desc), B, _ptr(params), _ptr(x_in), _ptr(grad_y),
                                                   _ptr(reg_cotangent), _ptr(zw), _ptr(grads), _ptr(gx), _ptr(ws), n,
                                                   _stream()))
        else:
            grads.zero_()
    return grads, gx


def fp32_fma_peak_tflops(iters: int = 4096, repeats: int = 5) -> float:
    """Measured FP32 FMA roofline of the current device: best of `repeats` timed burns."""
    lib = _lib.load()
    sink = torch.zeros(1, dtype=torch.float32, device="cuda")
    flops = ctypes.c_double(0.0)
    best = 0.0
    for it in range(repeats + 1):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _lib.check(lib.gpflow_fp32_fma_burn(iters, _ptr(sink), ctypes.byref(flops), _stream()))
        e1.record()
        e1.synchronize()
        if it > 0:
            best = max(best, flops.value / (e0.elapsed_time(e1) * 1e-3) / 1e12)
    return best


def tf32_mma_peak_tflops(iters: int = 20000, repeats: int = 3) -> float:
    """Measured dense TF32 rate of mma.sync.m16n8k8 on the current device (the instruction behind the generic kernel's
    3xTF32 contractions); the fp32-equivalent roof of a 3xTF32 contraction is a third of it."""
    lib = _lib.load()
    sink = torch.zeros(1, dtype=torch.float32, device="cuda")
    flops = ctypes.c_double(0.0)
    best = 0.0
    for it in range(repeats + 1):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _lib.check(lib.gpflow_tf32_mma_burn(iters, _ptr(sink), ctypes.byref(flops), _stream()))
        e1.record()
        e1.synchronize()
        if it > 0:
            best = max(best, flops.value / (e0.elapsed_time(e1) * 1e-3) / 1e12)
    return best
