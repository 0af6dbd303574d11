"""The chunked tensor-core path (gpflow_tc*.cuh: tcgen05 / TMEM / TMA GEMMs over slabs of samples) through the C ABI,
against the fp64 oracle on identical inputs, the generic kernel, and -- at the MEASURED hyper-parameters of bench.py
(N = 15 RK4 steps, weights N(0, w_std^2), the bench's synthetic batch) -- against vectors from the unmodified reference
(tests/golden/measured_*.npz, oracle/make_golden_measured.py).  Tolerance: 1e-5 relative, norm-wise (BASELINE.json)."""
import os

import numpy as np
import pytest
import torch

import gpflow_oracle as O
from golden_util import GOLDEN_DIR, rel

pytestmark = pytest.mark.gpu

SHAPES = [
    # d, nb, hidden, B, n_steps, weight std
    (5, 3, (64, 64), 300, 4, 0.05),          # aligned widths, ragged batch (slab padded to 512)
    (6, 6, (40,), 130, 4, 0.05),             # one hidden layer (no hidden GEMM), nb = d
    (10, 9, (50, 50, 50), 200, 4, 0.03),     # C3 shape: three layers, widths zero-padded to 64
    (17, 16, (100, 70), 260, 3, 0.02),       # 256-wide tile instantiation, widths padded to 128
    (3, 1, (33, 200), 129, 3, 0.06),         # nb = 1, narrow -> wide
    (12, 5, (128, 64), 257, 3, 0.03),        # wide -> narrow (Ub_1 does not alias T_2)
    (18, 17, (48,), 140, 3, 0.02),           # 64-wide last layer with nb, d > 16: several row tiles per group in tc_w3_small_kernel
    (4, 2, (40, 40, 40, 40), 300, 3, 0.05),   # four 64-wide layers: the value-row chain kernel at its largest shared-memory size
    (16, 5, (48, 48), 200, 3, 0.03),          # d = 16: the upper bound of the d-bounded narrow-net kernels (chain, first-layer gradient)
]


def _inputs(d, nb, hidden, B, N, std):
    spec = O.FlowSpec(d=d, nb=nb, hidden=list(hidden), n_steps=N)
    params = [p.float().double() for p in O.init_params(spec, std=std, seed=11 + d, dtype=torch.float64)]
    gen = torch.Generator().manual_seed(100 + d * nb)
    x_map = torch.randn(B, d, generator=gen, dtype=torch.float64).clamp(-2.5, 2.5).float().double()
    x_data = (1.1 * x_map.flip(-1) + 0.1).float().double()
    return spec, params, x_data, x_map


@pytest.mark.parametrize("d,nb,hidden,B,N,std", SHAPES)
def test_tc_train_step_vs_oracle(d, nb, hidden, B, N, std):
    import gpflow_b200 as G
    spec, params, x_data, x_map = _inputs(d, nb, hidden, B, N, std)
    r = O.train_step(params, x_data, x_map, spec, with_lagrangian=True, need_grad_x=True)
    cfg = G.FlowConfig(d=d, nb=nb, hidden=tuple(hidden), n_steps=N)
    assert "gpflow_tc" in G.kernel_variant(cfg, force_tc=True)
    flat = O.flatten_params(params).float().cuda()
    res = G.fused_train_step(cfg, flat, x_data.float().cuda(), x_map.float().cuda(), want_y=True, want_grad_x=True, force_tc=True)
    sc = res["scalars"].cpu().double()
    assert rel(res["y"].cpu(), r["y"]) < 1e-5, ("y", rel(res["y"].cpu(), r["y"]))
    assert abs(sc[0] - r["ot"]) < 1e-5 * abs(r["ot"]), ("ot", float(sc[0]), float(r["ot"]))
    assert float(sc[1]) == 0.0
    assert rel(res["grads"].cpu(), r["grads"]) < 1e-5, ("grads", rel(res["grads"].cpu(), r["grads"]))
    assert rel(res["grad_x"].cpu(), r["grad_x"]) < 2e-5, ("grad_x", rel(res["grad_x"].cpu(), r["grad_x"]))
    # run-to-run bit reproducibility (fixed reduction orders, one CTA per gradient partial)
    res2 = G.fused_train_step(cfg, flat, x_data.float().cuda(), x_map.float().cuda(), force_tc=True)
    assert torch.equal(res2["buf"], res["buf"])


@pytest.mark.parametrize("d,nb,hidden,B,N,std", SHAPES[:4])
def test_tc_forward_features_vs_generic_and_oracle(d, nb, hidden, B, N, std):
    """forward op: log-det, trajectory, Lagrangian, reverse integration, n_steps_eval; velocity call; backward op."""
    import gpflow_b200 as G
    from gpflow_b200 import ops_aux
    spec, params, x_data, x_map = _inputs(d, nb, hidden, B, N, std)
    cfg = G.FlowConfig(d=d, nb=nb, hidden=tuple(hidden), n_steps=N, n_steps_eval=N + 2)
    flat = O.flatten_params(params).float().cuda()
    xm = x_map.float().cuda()
    for reverse in (False, True):
        a = G.flow_forward(cfg, flat, xm, wrap=True, reverse=reverse, lagrangian=True, want_logdet=True, want_traj=True, force_tc=True)
        b = G.flow_forward(cfg, flat, xm, wrap=True, reverse=reverse, lagrangian=True, want_logdet=True, want_traj=True,
                           force_generic=True)
        assert rel(a["y"].cpu(), b["y"].cpu()) < 2e-6 and rel(a["traj"].cpu(), b["traj"].cpu()) < 2e-6
        assert rel(a["logdet"].cpu(), b["logdet"].cpu()) < 1e-4
        assert abs(float(a["scalars"][2]) - float(b["scalars"][2])) < 1e-5 * abs(float(b["scalars"][2]))
    # raw (cube-space) forward against the oracle
    g0 = torch.erf(x_map / 2 ** 0.5)
    raw = G.flow_forward(cfg, flat, g0.float().cuda(), wrap=False, training=True, force_tc=True)
    ref = O.flow_forward(params, g0.float().double(), spec, training=True)
    assert rel(raw["y"].cpu(), ref.out) < 2e-6
    # VelocityField.__call__
    flow = G.get_gp_flow(d, {"nb_neurons": list(hidden), "nb_div_blocks": nb}, N).cuda()
    with torch.no_grad():
        off = 0
        for p in flow.parameters():
            p.copy_(flat[off:off + p.numel()].view_as(p)); off += p.numel()
    v = ops_aux.velocity(flow.v, g0.float().cuda(), 0.3, force_tc=True)
    v_ref = O.velocity(params, g0.float().double(), torch.tensor(0.3, dtype=torch.float64), nb, O.eps_of(torch.float32))
    assert rel(v.cpu(), v_ref) < 1e-5
    # backward op with an arbitrary cotangent
    gy = torch.randn(B, d, generator=torch.Generator().manual_seed(5)).cuda()
    ga, gxa = G.flow_backward(cfg, flat, xm, gy, wrap=True, training=True, force_tc=True)
    gb, gxb = G.flow_backward(cfg, flat, xm, gy, wrap=True, training=True, force_generic=True)
    assert rel(ga.cpu(), gb.cpu()) < 5e-6 and rel(gxa.cpu(), gxb.cpu()) < 5e-6


def test_tc_accumulates_across_slabs(monkeypatch):
    """GPFLOW_TC_SLAB forces 256-sample slabs: gradients / scalars accumulate over three slabs, the last one ragged."""
    import gpflow_b200 as G
    d, nb, hidden, B, N, std = 10, 9, (50, 50, 50), 700, 3, 0.03
    spec, params, x_data, x_map = _inputs(d, nb, hidden, B, N, std)
    r = O.train_step(params, x_data, x_map, spec)
    cfg = G.FlowConfig(d=d, nb=nb, hidden=tuple(hidden), n_steps=N)
    flat = O.flatten_params(params).float().cuda()
    monkeypatch.setenv("GPFLOW_TC_SLAB", "256")
    res = G.fused_train_step(cfg, flat, x_data.float().cuda(), x_map.float().cuda(), want_y=True, force_tc=True)
    assert rel(res["y"].cpu(), r["y"]) < 1e-5
    assert abs(float(res["scalars"][0]) - float(r["ot"])) < 1e-5 * float(r["ot"])
    assert rel(res["grads"].cpu(), r["grads"]) < 1e-5, rel(res["grads"].cpu(), r["grads"])


def test_tc_is_the_default_for_large_wide_batches():
    import gpflow_b200 as G
    cfg = G.FlowConfig(d=10, nb=9, hidden=(50, 50, 50), n_steps=15)
    assert "gpflow_tc" in G.kernel_variant(cfg)                       # fp32, no Euler
    assert "generic" in G.kernel_variant(cfg, euler_coef=5e-5)        # Euler penalization stays on the generic kernel
    assert "generic" in G.kernel_variant(cfg, torch.float64)
    assert "sf2" in G.kernel_variant(G.FlowConfig(d=2, nb=1, hidden=(15, 15)))
    spec, params, x_data, x_map = _inputs(10, 9, (50, 50, 50), 6144, 2, 0.03)
    flat = O.flatten_params(params).float().cuda()
    cfg2 = G.FlowConfig(d=10, nb=9, hidden=(50, 50, 50), n_steps=2)
    a = G.fused_train_step(cfg2, flat, x_data.float().cuda(), x_map.float().cuda())                      # default: tc from B = 2400 for this shape
    b = G.fused_train_step(cfg2, flat, x_data.float().cuda(), x_map.float().cuda(), force_tc=True)
    c = G.fused_train_step(cfg2, flat, x_data.float().cuda(), x_map.float().cuda(), force_generic=True)
    assert torch.equal(a["buf"], b["buf"]) and rel(a["grads"].cpu(), c["grads"].cpu()) < 5e-6


def _measured(name):
    z = np.load(os.path.join(GOLDEN_DIR, name + "_f32.npz"))
    g = {k: z[k] for k in z.files}
    spec = O.FlowSpec(d=int(g["d"]), nb=int(g["nb"]), hidden=[int(h) for h in g["hidden"]], n_steps=int(g["N"]))
    params = O.init_params(spec, float(g["w_std"]), int(g["w_seed"]), torch.float32)
    return g, spec, params


@pytest.mark.parametrize("name", ["measured_c3_d10_nb9", "measured_c5_d64_nb63"])
@pytest.mark.parametrize("path", ["tc", "generic"])
def test_measured_configuration_vs_unmodified_reference(name, path):
    """bench.py's hyper-parameters (N = 15, weights N(0, 0.05^2) / N(0, 0.02^2), its synthetic batch): the reference's own
    y, log-det, OT cost, Lagrangian and gradients.  The reference's fp32 run is itself ~1e-6 from the fp64 truth."""
    import gpflow_b200 as G
    g, spec, params = _measured(name)
    cfg = G.FlowConfig(d=spec.d, nb=spec.nb, hidden=tuple(spec.hidden), n_steps=spec.n_steps)
    flat = O.flatten_params(params).cuda()
    kw = dict(force_tc=True) if path == "tc" else dict(force_generic=True)
    res = G.fused_train_step(cfg, flat, torch.from_numpy(g["x_data"]).cuda(), torch.from_numpy(g["x_map"]).cuda(),
                             lagrangian=(path == "generic"), want_y=True, want_grad_x=True, **kw)
    grads = res["grads"].cpu().double()
    assert rel(res["y"].cpu(), g["y"]) < 1e-5, rel(res["y"].cpu(), g["y"])
    assert abs(float(res["scalars"][0]) - float(g["ot"])) < 1e-5 * float(g["ot"])
    if path == "generic":
        assert abs(float(res["scalars"][2]) - float(g["lagrangian"])) < 1e-5 * float(g["lagrangian"])
    assert abs(float(grads.norm()) - float(g["grads_norm"])) < 1e-5 * float(g["grads_norm"])
    if "grads" in g:
        assert rel(grads, g["grads"]) < 1e-5, rel(grads, g["grads"])
    else:
        assert rel(grads[g["grad_idx"]], g["grads_sub"]) < 1e-5
        gen = np.random.default_rng(0)
        for k in range(4):
            proj = float((grads * torch.from_numpy(gen.standard_normal(grads.numel()))).sum())
            assert abs(proj - g["grads_proj"][k]) < 2e-5 * float(g["grads_norm"]) * np.sqrt(grads.numel()) / 30, (k, proj)
    assert rel(res["grad_x"].cpu(), g["grad_x"]) < 2e-5


@pytest.mark.parametrize("name,B", [("measured_c5_d64_nb63", 8), ("measured_c3_d10_nb9", 64)])
def test_measured_configuration_vs_fp64_oracle(name, B):
    """Same hyper-parameters, B samples of the bench's batch, fp64 oracle truth at 1e-5 (y, OT, gradients) through the
    tensor-core path: 60 evaluations of 3xTF32 GEMMs, segment-accumulated weight gradients."""
    import gpflow_b200 as G
    from bench import synth_batch
    g, spec, params = _measured(name)
    x_data, x_map = synth_batch(B, spec.d, 1234, "cpu")
    r = O.train_step([p.double() for p in params], x_data.double(), x_map.double(), spec, need_grad_x=True)
    cfg = G.FlowConfig(d=spec.d, nb=spec.nb, hidden=tuple(spec.hidden), n_steps=spec.n_steps)
    res = G.fused_train_step(cfg, O.flatten_params(params).cuda(), x_data.cuda(), x_map.cuda(), want_y=True, want_grad_x=True,
                             force_tc=True)
    assert rel(res["y"].cpu(), r["y"]) < 1e-5, rel(res["y"].cpu(), r["y"])
    assert abs(float(res["scalars"][0]) - float(r["ot"])) < 1e-5 * float(r["ot"])
    assert rel(res["grads"].cpu(), r["grads"]) < 1e-5, rel(res["grads"].cpu(), r["grads"])
    assert rel(res["grad_x"].cpu(), r["grad_x"]) < 2e-5


@pytest.mark.parametrize("d,nb,hidden,w_std,n_sub", [(10, 9, (50, 50, 50), 0.05, 32), (64, 63, (256, 256), 0.02, 4)])
def test_bench_rows_at_full_size_2p14_properties(d, nb, hidden, w_std, n_sub):
    """The C3 / C5 rows of bench.py at their full per-GPU size (2^14 samples, N = 15, bench weights), through size-independent
    properties: (i) the fused [grads | ot] buffer of the batch equals the SUM of the buffers of its two halves computed with
    loss_scale = 1/2^14 (what the data-parallel all-reduce computes; different slab counts / split sizes on both sides);
    (ii) y of a strided subsample equals the fp64 oracle on those samples; (iii) the OT scalar equals the mean over the
    transported points the kernels wrote; (iv) a second launch is bit-identical."""
    import gpflow_b200 as G
    B, N = 1 << 14, 15
    spec = O.FlowSpec(d=d, nb=nb, hidden=list(hidden), n_steps=N)
    params = O.init_params(spec, w_std, 0, torch.float32)
    flat = O.flatten_params(params).cuda()
    g = torch.Generator(device="cuda").manual_seed(1234)
    x_map = torch.randn(B, d, generator=g, device="cuda")
    x_data = 0.9 * x_map.roll(1, dims=1) + 0.1
    cfg = G.FlowConfig(d=d, nb=nb, hidden=tuple(hidden), n_steps=N)
    assert "gpflow_tc" in G.kernel_variant(cfg)
    full = G.fused_train_step(cfg, flat, x_data, x_map, want_y=True)
    again = G.fused_train_step(cfg, flat, x_data, x_map)["buf"].clone()
    assert torch.equal(again, full["buf"])
    acc = torch.zeros_like(full["buf"], dtype=torch.float64)
    for r in range(2):
        lo, hi = r * (B // 2), (r + 1) * (B // 2)
        acc += G.fused_train_step(cfg, flat, x_data[lo:hi], x_map[lo:hi], loss_scale=1.0 / B)["buf"].double()
    assert rel(acc.cpu(), full["buf"].double().cpu()) < 5e-6, rel(acc.cpu(), full["buf"].double().cpu())
    idx = torch.arange(0, B, B // n_sub, device="cuda")
    r64 = O.train_step([p.double() for p in params], x_data[idx].cpu().double(), x_map[idx].cpu().double(), spec)
    assert rel(full["y"][idx].cpu(), r64["y"]) < 1e-5
    ot = ((x_data.double() - full["y"].double()) ** 2).sum(1).mean()
    assert abs(float(full["scalars"][0]) - float(ot)) < 2e-6 * float(ot)


def test_tc_train_step_captured_in_a_cuda_graph():
    """The tensor-core path is ~2 200 launches per step: a caller may capture the call once (after a warm-up call that sets
    the kernel attributes) and replay it.  Replays are bit-identical to the eager call and follow new parameter values."""
    import gpflow_b200 as G
    d, nb, hidden, B, N, std = 10, 9, (50, 50, 50), 3072, 3, 0.03
    spec, params, x_data, x_map = _inputs(d, nb, hidden, B, N, std)
    cfg = G.FlowConfig(d=d, nb=nb, hidden=tuple(hidden), n_steps=N)
    flat = O.flatten_params(params).float().cuda()
    xd, xm = x_data.float().cuda(), x_map.float().cuda()
    out = torch.zeros(cfg.n_params + 3, device="cuda")
    assert "gpflow_tc" in G.kernel_variant(cfg)
    want = G.fused_train_step(cfg, flat, xd, xm, out=out)["buf"].clone()          # warm-up outside the capture
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    out.zero_()
    with torch.cuda.graph(graph):
        G.fused_train_step(cfg, flat, xd, xm, out=out)
    out.zero_()
    graph.replay()
    torch.cuda.synchronize()
    assert torch.equal(out, want)
    flat.mul_(1.1)
    want2 = G.fused_train_step(cfg, flat, xd, xm)["buf"].clone()
    graph.replay()
    torch.cuda.synchronize()
    assert torch.equal(out, want2)


def test_module_path_replays_a_captured_graph_for_narrow_nets(monkeypatch):
    """GPFlowNetwork.fused_training_step with the fused optimiser's stable buffers: the tensor-core train step of a narrow
    net is captured once and replayed (network.py: _graphed_train_step).  Same losses and parameters, bit for bit, as with
    eager launches (GPFLOW_TC_GRAPH=0); the replays are counted as launches of the path."""
    import gpflow_b200 as G
    from gpflow_b200 import ops

    def run(graph_on):
        monkeypatch.setenv("GPFLOW_TC_GRAPH", "1" if graph_on else "0")
        torch.manual_seed(0)
        flow = G.get_gp_flow(10, {"nb_neurons": [50, 50, 50], "nb_div_blocks": 9}, 3).cuda()
        with torch.no_grad():
            for p in flow.parameters():
                p.normal_(0.0, 0.03)
        net = G.GPFlowNetwork(flow, map_module=None)
        opt = net.configure_fused_optimizer(lr=2e-3, weight_decay=1e-4)
        g = torch.Generator(device="cuda").manual_seed(3)
        losses = []
        n0 = ops.tc_launches_total()
        for i in range(4):
            x_map = torch.randn(3072, 10, generator=g, device="cuda")
            x_data = 0.9 * x_map.roll(1, 1) + 0.1
            loss, _ = net.fused_training_step(x_data, x_map)
            opt.step()
            losses.append(loss.clone())
        torch.cuda.synchronize()
        return torch.stack(losses), opt.flat.clone(), ops.tc_launches_total() - n0

    la, pa, na = run(True)
    lb, pb, nb_ = run(False)
    assert torch.equal(la, lb) and torch.equal(pa, pb)
    assert na >= nb_ > 0            # warm-up + capture + 4 replays against 4 eager steps
