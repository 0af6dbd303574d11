"""Auxiliary GPU evaluations used by the module mirrors outside the train step (plots / trajectories)."""
import ctypes

import torch

from . import _lib, ops


def velocity(vf, x: torch.Tensor, t, force_tc=False) -> torch.Tensor:
    """VelocityField.__call__(x, t) (velocity_field.py:118-126) on the GPU."""
    cfg = ops.FlowConfig(d=vf.dim, nb=vf.nb_div_blocks, hidden=tuple(vf.v.neurons[1:-1]), n_steps=1)
    x = ops._check_cuda("x", x)
    flat = vf.flat_parameters().detach()
    desc = ops.make_desc(cfg, x.dtype, force_generic=not force_tc, force_tc=force_tc)
    out = torch.empty_like(x)
    lib = _lib.load()
    with torch.cuda.device(x.device):
        ws, n = ops._workspace(desc, x.shape[0], _lib.OP_FORWARD, x.device)
        _lib.check(lib.gpflow_velocity(ctypes.byref(desc), x.shape[0], ops._ptr(flat), ops._ptr(x),
                                       ctypes.c_double(float(t)), ops._ptr(out), ops._ptr(ws), n, ops._stream()))
    return out


def velocity_static(cfg, flat_params: torch.Tensor, x: torch.Tensor, t, out: torch.Tensor) -> torch.Tensor:
    """`gpflow_velocity_static` (include/gpflow_b200.h; no counterpart in the reference): v(x, t) into the caller's `out`
    with the library's built-in 32 MiB scratch instead of a workspace -- this wrapper and the call allocate nothing
    (all tensors must already be contiguous CUDA tensors of one dtype).  Calls that overlap on the GPU race on the scratch.
    Raises (GPFLOW_ERR_STATIC_SCRATCH) when the shape / batch needs more than the scratch holds."""
    for name, ten in (("flat_params", flat_params), ("x", x), ("out", out)):
        if not ten.is_cuda or not ten.is_contiguous() or ten.dtype != x.dtype:
            raise RuntimeError("gpflow_b200: `%s` must be a contiguous CUDA tensor of x's dtype" % name)
    if out.shape != x.shape:
        raise ValueError("gpflow_b200: `out` must have x's shape")
    desc = ops.make_desc(cfg, x.dtype, force_generic=True)
    with torch.cuda.device(x.device):
        _lib.check(_lib.load().gpflow_velocity_static(ctypes.byref(desc), x.shape[0], ops._ptr(flat_params), ops._ptr(x),
                                                     ctypes.c_double(float(t)), ops._ptr(out), ops._stream()))
    return out


def potentials(vfield, x, want_f=True, want_jac=True):
    """VectorField.eval / .jacobian (vector_field.py:75-128) on the GPU: (f [B, dim_out], J [B, dim_out, dim_in]).
    `x` has the network's full input width (the velocity field feeds it [x; t]).  Not differentiable: the fused flow ops
    carry the gradients of the path."""
    x2 = ops._check_cuda("x", x if x.dim() == 2 else x.reshape(1, -1))
    widths = (ctypes.c_int32 * len(vfield.neurons))(*vfield.neurons)
    if x2.shape[1] != vfield.neurons[0]:
        raise ValueError("gpflow_b200: x has %d columns, the network takes %d" % (x2.shape[1], vfield.neurons[0]))
    flat = vfield.flat_parameters().detach().to(x2.dtype)
    B = x2.shape[0]
    f = torch.empty(B, vfield.neurons[-1], dtype=x2.dtype, device=x2.device) if want_f else None
    jac = torch.empty(B, vfield.neurons[-1], vfield.neurons[0], dtype=x2.dtype, device=x2.device) if want_jac else None
    with torch.cuda.device(x2.device):
        _lib.check(_lib.load().gpflow_potentials(ops._dtype_code(x2.dtype), len(vfield.neurons) - 1, widths, B, ops._ptr(flat),
                                                 ops._ptr(x2), ops._ptr(f), ops._ptr(jac), ops._stream()))
    if x.dim() == 1:
        f, jac = (None if f is None else f[0]), (None if jac is None else jac[0])
    return f, jac
