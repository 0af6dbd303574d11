"""The multi-GPU hook of the C ABI (include/gpflow_b200.h: gpflow_comm_unique_id / gpflow_comm_create / gpflow_allreduce_grads /
gpflow_comm_destroy; SURVEY.md 8b "Multi-GPU hook", 8e).  CPU: argument checking and the run-time NCCL binding.  GPU (one
device): a one-rank communicator -- the in-place SUM over one rank is the identity -- on the fused buffer of a real train
step, fp32 and fp64, on a side stream.  The two-rank comparison with torch.distributed is tools/native_allreduce_check.py
(run under torchrun on 2 GPUs; its output is profiles/r2_native_allreduce_n2.txt)."""
import ctypes

import pytest
import torch


def _lib():
    from gpflow_b200 import build, _lib
    build.build()
    return _lib, _lib.load()


def test_argument_checks_and_error_string():
    L, lib = _lib()
    assert lib.gpflow_comm_unique_id(None) == L.ERR_NULL_POINTER
    assert lib.gpflow_comm_create(None, 1, 0, None) == L.ERR_NULL_POINTER
    ident = (ctypes.c_char * 128)()
    comm = ctypes.c_void_p(0)
    assert lib.gpflow_comm_create(ctypes.cast(ident, ctypes.c_void_p), 0, 0, ctypes.byref(comm)) == L.ERR_BAD_DESC
    assert lib.gpflow_comm_create(ctypes.cast(ident, ctypes.c_void_p), 2, 2, ctypes.byref(comm)) == L.ERR_BAD_DESC
    assert lib.gpflow_allreduce_grads(None, 4, L.F32, None, None) == L.ERR_NULL_POINTER
    assert lib.gpflow_comm_destroy(None) == 0
    assert b"NCCL" in lib.gpflow_error_string(L.ERR_COLLECTIVE)


def test_nccl_is_bound_at_run_time_and_yields_an_id():
    """ncclGetUniqueId needs no device: inside a PyTorch process the already loaded NCCL is found."""
    L, lib = _lib()
    ident = (ctypes.c_char * 128)()
    assert lib.gpflow_comm_unique_id(ctypes.cast(ident, ctypes.c_void_p)) == 0
    assert any(b != b"\x00" for b in ident)


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_one_rank_communicator_on_the_train_step_buffer(dtype):
    import gpflow_oracle as O
    import gpflow_b200 as G
    from gpflow_b200.data_parallel import NativeAllReduce
    spec = O.FlowSpec(d=2, nb=1, hidden=[15, 15], n_steps=6)
    flat = O.flatten_params(O.init_params(spec, 0.3, 1, dtype)).cuda()
    g = torch.Generator().manual_seed(2)
    x_map = torch.randn(1000, 2, generator=g, dtype=dtype).cuda()
    x_data = 1.2 * x_map.flip(-1) + 0.1
    cfg = G.FlowConfig(d=2, nb=1, hidden=(15, 15), n_steps=6)
    ar = NativeAllReduce("cuda:0")
    assert ar.world == 1
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        buf = G.fused_train_step(cfg, flat, x_data, x_map)["buf"]
        want = buf.clone()
        ar.allreduce_(buf)                       # enqueued behind the train step on the same (side) stream
    side.synchronize()
    assert torch.equal(buf, want) and float(buf.abs().sum()) > 0
    with pytest.raises(ValueError):
        ar.allreduce_(buf.cpu())
    ar.close()
