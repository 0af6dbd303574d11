"""Frozen BNAF map forward: oracle restatement and CUDA kernel against golden vectors produced
by the unmodified reference modules of pretrained_models/{eight_gaussians,moons}/bnaf (oracle/make_golden_bnaf.py)."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import bnaf_oracle as BO  # noqa: E402

DATA = ["eight_gaussians", "moons"]


def _golden(name):
    z = np.load(os.path.join(ROOT, "tests", "golden", "bnaf_%s_f32.npz" % name))
    g = {k: z[k] for k in z.files}
    return g, int(g["dim"]), int(g["n_neurons"]), int(g["n_layers"])


def _rel(a, b):
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


@pytest.mark.parametrize("name", DATA)
def test_oracle_matches_reference_golden(name):
    g, dim, nn, nl = _golden(name)
    flat, x = torch.from_numpy(g["params"]), torch.from_numpy(g["x"])
    assert flat.numel() == sum(o * i + 2 * o for o, i in BO.layer_shapes(dim, nn, nl)) == 2372
    y32, ld32 = BO.bnaf_forward(flat, x, dim, nn, nl)
    assert _rel(y32, torch.from_numpy(g["y"])) < 2e-6
    assert _rel(ld32, torch.from_numpy(g["logdet"])) < 2e-6
    y64, ld64 = BO.bnaf_forward(flat.double(), x.double(), dim, nn, nl)
    assert _rel(y64.float(), torch.from_numpy(g["y"])) < 2e-6
    assert _rel(ld64.float(), torch.from_numpy(g["logdet"])) < 2e-6


def test_oracle_logdet_is_the_jacobian_determinant():
    """the block logsumexp equals log|det dy/dx| of the map (autograd Jacobian, fp64)"""
    g, dim, nn, nl = _golden("moons")
    flat = torch.from_numpy(g["params"]).double()
    x = torch.from_numpy(g["x"][:8]).double()
    _, ld = BO.bnaf_forward(flat, x, dim, nn, nl)
    for b in range(8):
        J = torch.autograd.functional.jacobian(lambda v: BO.bnaf_forward(flat, v[None], dim, nn, nl)[0][0], x[b])
        assert abs(float(torch.linalg.slogdet(J)[1]) - float(ld[b])) < 1e-9


def _need_cuda():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")


@pytest.mark.gpu
@pytest.mark.parametrize("name", DATA)
def test_cuda_matches_reference_golden(name):
    _need_cuda()
    import gpflow_b200 as G
    g, dim, nn, nl = _golden(name)
    m = G.BNAFMap(dim, nn, nl, torch.from_numpy(g["params"])).cuda()
    y, ld = m(torch.from_numpy(g["x"]).cuda())
    assert _rel(y.cpu(), torch.from_numpy(g["y"])) < 1e-5
    assert _rel(ld.cpu(), torch.from_numpy(g["logdet"])) < 1e-5


@pytest.mark.gpu
@pytest.mark.parametrize("B", [1, 31, 255, 257, 4099, 200000])
def test_cuda_matches_fp64_oracle_ragged(B):
    _need_cuda()
    import gpflow_b200 as G
    g, dim, nn, nl = _golden("eight_gaussians")
    flat = torch.from_numpy(g["params"])
    m = G.BNAFMap(dim, nn, nl, flat).cuda()
    gen = torch.Generator().manual_seed(B)
    x = 2.0 * torch.randn(B, 2, generator=gen)
    y, ld = m(x.cuda())
    yo, ldo = BO.bnaf_forward(flat.double(), x.double(), dim, nn, nl)
    assert _rel(y.cpu().double(), yo) < 1e-5
    assert _rel(ld.cpu().double(), ldo) < 1e-5
    assert float((y.cpu().double() - yo).abs().max()) < 2e-5 * max(1.0, float(yo.abs().max()))


@pytest.mark.gpu
def test_cuda_empty_batch_and_unsupported_shapes():
    _need_cuda()
    import gpflow_b200 as G
    g, dim, nn, nl = _golden("moons")
    m = G.BNAFMap(dim, nn, nl, torch.from_numpy(g["params"])).cuda()
    y, ld = m(torch.empty(0, 2, device="cuda"))
    assert y.shape == (0, 2) and ld.shape == (0,)
    with pytest.raises(NotImplementedError):
        G.BNAFMap(3, 48, 4, torch.zeros(10))
    with pytest.raises(RuntimeError):
        m(torch.zeros(4, 2))            # CPU tensor: there is no CPU path


@pytest.mark.gpu
def test_map_in_front_of_the_train_step():
    """x -> BNAFMap -> fused train step equals feeding the reference's x_map (C1 golden trajectory, step 0)."""
    _need_cuda()
    import gpflow_b200 as G
    z = np.load(os.path.join(ROOT, "tests", "golden", "train_c1_eight_gaussians_f32.npz"))
    g, dim, nn, nl = _golden("eight_gaussians")
    m = G.BNAFMap(dim, nn, nl, torch.from_numpy(g["params"])).cuda()
    x = torch.from_numpy(z["x"][0]).cuda()
    x_map_ref = torch.from_numpy(z["x_map"][0])
    x_map, _ = m(x)
    assert _rel(x_map.cpu(), x_map_ref) < 1e-5
    flow = G.get_gp_flow(2, {"nb_neurons": [15, 15]}, 15).cuda()
    off = 0
    p0 = torch.from_numpy(z["params0"]).cuda()
    with torch.no_grad():
        for p in flow.parameters():
            p.copy_(p0[off:off + p.numel()].view_as(p))
            off += p.numel()
    net = G.GPFlowNetwork(flow, map_module=m)
    loss, _ = net.fused_training_step(x)        # x_map = self.map(x)[0] inside, network.py:76-78
    assert abs(float(loss) - float(z["ot"][0])) < 2e-5 * abs(float(z["ot"][0]))
