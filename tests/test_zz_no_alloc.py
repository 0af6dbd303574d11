"""`gpflow_velocity_static` (include/gpflow_b200.h): the velocity evaluation with the library's file-scope device scratch
instead of a caller workspace, for callers that cannot allocate.  It has no counterpart in the reference,
so the check is against the existing entry point `gpflow_velocity` -- same inputs through
the C-ABI, two calls in a row with different inputs (the scratch keeps the previous call's contents) -- and against the
fp64 oracle of `VelocityField.__call__` (velocity_field.py:118-126).  Collected last on purpose;
passes alone in a fresh process and after the other tests."""
import ctypes

import pytest
import torch

import gpflow_oracle as O
from golden_util import rel

pytestmark = pytest.mark.gpu

SHAPES = [  # (d, nb, hidden, B, dtype)
    (2, 1, (15, 15), 4096, torch.float32),       # on-chip scratch placement, any B fits
    (2, 1, (15, 15), 1000, torch.float64),
    (5, 3, (5, 5, 5), 777, torch.float32),       # ragged batch
    (10, 9, (50, 50, 50), 256, torch.float32),   # per-warp scratch in the device array (22.5 of the 32 MiB)
    (10, 9, (50, 50, 50), 100, torch.float64),
]


def _setup(d, nb, hidden, dtype, seed):
    from gpflow_b200 import ops
    cfg = ops.FlowConfig(d=d, nb=nb, hidden=tuple(hidden), n_steps=1)
    spec = O.FlowSpec(d=d, nb=nb, hidden=list(hidden), n_steps=1)
    gen = torch.Generator().manual_seed(seed)
    params = O.init_params(spec, 0.6, seed, dtype=dtype)       # weights and biases ~ N(0, 0.6^2), rounded to the dtype
    flat = O.flatten_params(params).cuda()
    return cfg, [p.double() for p in params], flat, gen


def _existing(lib, cfg, flat, x, t):
    """The existing entry point, caller-provided workspace (what ops_aux.velocity does)."""
    from gpflow_b200 import _lib, ops
    desc = ops.make_desc(cfg, x.dtype, force_generic=True)
    n = ctypes.c_size_t()
    _lib.check(lib.gpflow_query_workspace(ctypes.byref(desc), x.shape[0], _lib.OP_FORWARD, ctypes.byref(n)))
    ws = torch.empty(n.value + 256, dtype=torch.uint8, device=x.device)
    out = torch.full_like(x, float("nan"))
    _lib.check(lib.gpflow_velocity(ctypes.byref(desc), x.shape[0], ops._ptr(flat), ops._ptr(x), ctypes.c_double(t),
                                   ops._ptr(out), ops._ptr(ws), n.value, ops._stream()))
    return out


def _static(lib, cfg, flat, x, t, out):
    from gpflow_b200 import ops
    desc = ops.make_desc(cfg, x.dtype, force_generic=True)
    return lib.gpflow_velocity_static(ctypes.byref(desc), x.shape[0], ops._ptr(flat), ops._ptr(x), ctypes.c_double(t),
                                      ops._ptr(out), ops._stream())


@pytest.mark.parametrize("d,nb,hidden,B,dtype", SHAPES)
def test_static_scratch_velocity_matches_existing_entry_point(d, nb, hidden, B, dtype):
    from gpflow_b200 import _lib
    lib = _lib.load()
    cfg, params, flat, gen = _setup(d, nb, hidden, dtype, seed=4200 + d)
    tol_entry = 1e-6 if dtype == torch.float32 else 1e-12      # same kernel, same arithmetic: expected identical
    tol_oracle = 1e-5 if dtype == torch.float32 else 1e-10     # the parity bar of the other tests
    out = torch.empty(B, d, dtype=dtype, device="cuda")
    free_before = None
    for call, t in enumerate((0.3, 0.85)):                     # two calls in a row, different points and times
        x = (2 * torch.rand(B, d, generator=gen, dtype=torch.float64) - 1).mul(0.98).to(dtype).cuda()
        want = _existing(lib, cfg, flat, x, t)
        torch.cuda.synchronize()
        if call == 1:
            free_before = torch.cuda.mem_get_info()[0]
        out.fill_(float("nan"))
        assert _static(lib, cfg, flat, x, t, out) == 0
        torch.cuda.synchronize()
        if call == 1:                                          # (the first call may load the module: set-up)
            assert torch.cuda.mem_get_info()[0] == free_before, "the call changed the device's free memory"
        assert torch.isfinite(out).all()
        assert rel(out.cpu(), want.cpu()) <= tol_entry, (call, rel(out.cpu(), want.cpu()))
        v_ref = O.velocity(params, x.cpu().double(), torch.tensor(t, dtype=torch.float64), nb, O.eps_of(dtype))
        assert rel(out.cpu(), v_ref) < tol_oracle, (call, rel(out.cpu(), v_ref))


def test_static_scratch_limit_and_argument_errors():
    from gpflow_b200 import _lib, ops
    lib = _lib.load()
    cfg, _, flat, gen = _setup(10, 9, (50, 50, 50), torch.float32, seed=7)
    x = torch.zeros(1024, 10, device="cuda")
    out = torch.full_like(x, 7.0)
    rc = _static(lib, cfg, flat, x, 0.0, out)                  # 88.8 MB of scratch needed > 32 MiB
    assert rc == _lib.ERR_STATIC_SCRATCH
    assert b"GPFLOW_STATIC_SCRATCH_BYTES" in lib.gpflow_error_string(rc)
    torch.cuda.synchronize()
    assert bool((out == 7.0).all())                            # nothing was launched
    big = ops.FlowConfig(d=64, nb=63, hidden=(256, 256), n_steps=1)
    desc = ops.make_desc(big, torch.float32)
    assert lib.gpflow_velocity_static(ctypes.byref(desc), 1, ops._ptr(flat), ops._ptr(x), ctypes.c_double(0.0),
                                      ops._ptr(out), ops._stream()) == _lib.ERR_STATIC_SCRATCH
    # argument checks are the existing entry point's
    desc = ops.make_desc(cfg, torch.float32)
    assert lib.gpflow_velocity_static(None, 4, ops._ptr(flat), ops._ptr(x), ctypes.c_double(0.0), ops._ptr(out),
                                      ops._stream()) == 3      # GPFLOW_ERR_NULL_POINTER
    assert lib.gpflow_velocity_static(ctypes.byref(desc), 4, None, ops._ptr(x), ctypes.c_double(0.0), ops._ptr(out),
                                      ops._stream()) == 3
    assert lib.gpflow_velocity_static(ctypes.byref(desc), 0, ops._ptr(flat), ops._ptr(x), ctypes.c_double(0.0),
                                      ops._ptr(out), ops._stream()) == 0     # empty batch: nothing to do


def test_static_scratch_python_wrapper_and_graph_capture():
    """The use the entry point exists for: the call captured into a CUDA graph (where cudaMalloc is not allowed) and replayed."""
    from gpflow_b200 import _lib, ops_aux
    lib = _lib.load()
    cfg, params, flat, gen = _setup(5, 3, (5, 5, 5), torch.float32, seed=99)
    x = (2 * torch.rand(300, 5, generator=gen, dtype=torch.float64) - 1).mul(0.9).float().cuda()
    out = torch.zeros_like(x)
    ops_aux.velocity_static(cfg, flat, x, 0.5, out)            # set-up call on this device, outside the capture
    want = _existing(lib, cfg, flat, x, 0.5)
    torch.cuda.synchronize()
    assert rel(out.cpu(), want.cpu()) <= 1e-6
    graph = torch.cuda.CUDAGraph()
    out.zero_()
    with torch.cuda.graph(graph):
        ops_aux.velocity_static(cfg, flat, x, 0.5, out)
    x2 = (2 * torch.rand(300, 5, generator=gen, dtype=torch.float64) - 1).mul(0.9).float().cuda()
    x.copy_(x2)                                                # new inputs in the captured buffers
    graph.replay()
    torch.cuda.synchronize()
    want2 = _existing(lib, cfg, flat, x, 0.5)
    torch.cuda.synchronize()
    assert rel(out.cpu(), want2.cpu()) <= 1e-6
    with pytest.raises(RuntimeError, match="GPFLOW_STATIC_SCRATCH_BYTES"):
        big_cfg, _, big_flat, _ = _setup(10, 9, (50, 50, 50), torch.float32, seed=3)
        xb = torch.zeros(2048, 10, device="cuda")
        ops_aux.velocity_static(big_cfg, big_flat, xb, 0.0, torch.empty_like(xb))
