"""The packed forward weights of the 2-D kernels live in ONE constant-bank array per device, rewritten by every call
(gpflow_sf2r.cuh: sf2r_cw; launch_sf2r / sf2r_launch_guarded).  Calls with DIFFERENT parameters issued back to back on
different streams must still see their own weights: the launcher orders them on the device with an event recorded behind
the previous call's kernel.  Without that guard the second call's copy would overwrite the array while the first kernel is
reading it.  Also: the call captured in a CUDA graph and replayed."""
import pytest
import torch

import gpflow_oracle as O

pytestmark = pytest.mark.gpu

SPEC = O.FlowSpec(d=2, nb=1, hidden=[15, 15], n_steps=15)


def _cfg():
    from gpflow_b200 import FlowConfig
    return FlowConfig(d=2, nb=1, hidden=(15, 15), n_steps=15)


def test_interleaved_streams_with_different_parameters():
    import gpflow_b200 as G
    B = 1 << 16                                  # ~0.5 ms per call: the calls of the two streams would overlap
    g = torch.Generator().manual_seed(3)
    x_map = torch.randn(B, 2, generator=g).clamp(-3, 3).cuda()
    x_data = (1.3 * x_map.flip(-1) + 0.2)
    pa = O.flatten_params(O.init_params(SPEC, 0.3, 1, torch.float32)).cuda()
    pb = O.flatten_params(O.init_params(SPEC, 0.5, 2, torch.float32)).cuda()
    zw = torch.randn(15, 2, 2, generator=g).cuda()
    cfg = _cfg()
    # serial references (one stream, synchronised)
    ref = {}
    for name, p in (("a", pa), ("b", pb)):
        ref[name] = G.fused_train_step(cfg, p, x_data, x_map)["buf"].clone()
        ref[name + "e"] = G.fused_train_step(cfg, p, x_data, x_map, euler_coef=0.3, zw=zw)["buf"].clone()
        ref[name + "f"] = G.flow_forward(cfg, p, x_map, wrap=True)["y"].clone()
    torch.cuda.synchronize()
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    outs = []
    for it in range(6):
        with torch.cuda.stream(s1):
            ra = G.fused_train_step(cfg, pa, x_data, x_map)["buf"]
            rae = G.fused_train_step(cfg, pa, x_data, x_map, euler_coef=0.3, zw=zw)["buf"] if it % 2 else None
        with torch.cuda.stream(s2):
            rb = G.fused_train_step(cfg, pb, x_data, x_map)["buf"]
            rbf = G.flow_forward(cfg, pb, x_map, wrap=True)["y"]
        outs.append((ra, rae, rb, rbf))
    torch.cuda.synchronize()
    for ra, rae, rb, rbf in outs:
        assert torch.equal(ra, ref["a"]) and torch.equal(rb, ref["b"]) and torch.equal(rbf, ref["bf"])
        if rae is not None:
            assert torch.equal(rae, ref["ae"])


def test_train_step_captured_in_a_cuda_graph():
    import gpflow_b200 as G
    B = 4096
    g = torch.Generator().manual_seed(5)
    x_map = torch.randn(B, 2, generator=g).clamp(-3, 3).cuda()
    x_data = 1.2 * x_map.flip(-1) + 0.1
    p = O.flatten_params(O.init_params(SPEC, 0.3, 4, torch.float32)).cuda()
    cfg = _cfg()
    out = torch.zeros(cfg.n_params + 3, device="cuda")
    want = G.fused_train_step(cfg, p, x_data, x_map, out=out)["buf"].clone()      # warm-up outside the capture
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    out.zero_()
    with torch.cuda.graph(graph):
        G.fused_train_step(cfg, p, x_data, x_map, out=out)
    out.zero_()
    graph.replay()
    torch.cuda.synchronize()
    assert torch.equal(out, want)
    p2 = O.flatten_params(O.init_params(SPEC, 0.4, 6, torch.float32)).cuda()
    want2 = G.fused_train_step(cfg, p2, x_data, x_map)["buf"].clone()
    p.copy_(p2)                                   # new parameters in the captured buffer
    graph.replay()
    torch.cuda.synchronize()
    assert torch.equal(out, want2)
