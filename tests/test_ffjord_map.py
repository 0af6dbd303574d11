"""Frozen FFJORD map + Gaussian-noise data set: oracle restatement and CUDA kernel against golden
vectors produced by the unmodified reference modules of pretrained_models/dsprites/ffjord (oracle/make_golden_ffjord.py:
the reference's ODEfunc vector field / brute-force divergence, and its FFJORD -> SequentialFlow -> CNF path driven by a
fixed-grid classical RK4 in place of the missing torchdiffeq)."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import ffjord_oracle as FO  # noqa: E402


def _golden(prec):
    z = np.load(os.path.join(ROOT, "tests", "golden", "ffjord_dsprites_%s.npz" % prec))
    g = {k: z[k] for k in z.files}
    spec = FO.FfjordSpec(d=int(g["d"]), hidden=[int(h) for h in g["hidden"]], n_blocks=int(g["n_blocks"]),
                         n_steps=int(g["n_steps"]), t_end=[float(t) for t in g["t_end"]])
    return g, spec


def _rel(a, b):
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def _rand_params(spec, seed, dtype, scale=0.5):
    g = torch.Generator().manual_seed(seed)
    n = spec.dims()
    parts = []
    for _ in range(spec.n_blocks):
        for l in range(len(n) - 1):
            parts.append(torch.randn(n[l + 1] * n[l], generator=g, dtype=torch.float64) * (scale / n[l] ** 0.5))
            parts.append(torch.randn(4 * n[l + 1], generator=g, dtype=torch.float64) * 0.3)
    return torch.cat(parts).to(dtype)


# ---------------------------------------------------------------- CPU: the oracle against the reference's own outputs
@pytest.mark.parametrize("prec,tol", [("f32", 2e-6), ("f64", 1e-13)])
def test_oracle_vector_field_matches_reference_odefunc(prec, tol):
    g, spec = _golden(prec)
    flat, x = torch.from_numpy(g["params"]), torch.from_numpy(g["x"])
    assert flat.numel() == spec.param_count() == 17784
    blocks = FO.unpack(flat, spec)
    for b in range(spec.n_blocks):
        for ti, t in enumerate(g["vf_t"]):
            f, div = FO.odenet(float(t), x, blocks[b], want_div=True)
            assert _rel(f, torch.from_numpy(g["vf_f"][b, ti])) < tol
            assert _rel(div, torch.from_numpy(g["vf_div"][b, ti])) < 5 * tol


@pytest.mark.parametrize("prec,tol", [("f32", 5e-6), ("f64", 1e-12)])
def test_oracle_map_matches_reference_chain(prec, tol):
    g, spec = _golden(prec)
    flat, x = torch.from_numpy(g["params"]), torch.from_numpy(g["x"])
    z, ld = FO.integrate(flat, x, spec, want_logdet=True)
    assert _rel(z, torch.from_numpy(g["z"])) < tol
    assert _rel(ld, torch.from_numpy(g["logdet"])) < tol
    xb, _ = FO.integrate(flat, torch.from_numpy(g["z"]), spec, reverse=True)
    assert _rel(xb, torch.from_numpy(g["xb"])) < tol


def test_oracle_logdet_is_the_jacobian_determinant():
    """the integrated trace equals log|det dz/dx| of the discrete map up to the RK4 error: compare with the autograd
    Jacobian of the SAME discrete map at a step count where both agree (fp64)."""
    spec = FO.FfjordSpec(d=3, hidden=[6, 5], n_blocks=2, n_steps=48)
    flat = _rand_params(spec, 3, torch.float64)
    x = torch.randn(4, 3, generator=torch.Generator().manual_seed(1), dtype=torch.float64)
    _, ld = FO.integrate(flat, x, spec, want_logdet=True)
    for b in range(4):
        J = torch.autograd.functional.jacobian(lambda v: FO.integrate(flat, v[None], spec)[0][0], x[b])
        assert abs(float(torch.linalg.slogdet(J)[1]) - float(ld[b])) < 1e-7


def test_step_halving_error_recorded_in_golden():
    g, _ = _golden("f64")
    assert g["err_halving"][0] < 2e-3      # 16 steps per block on the wide test points (integrator error, not parity)


def test_abi_ffjord_desc_and_param_count_cpu():
    """no compute: the library loads without a GPU, the ctypes descriptor matches the C struct (a wrong layout would move
    n_blocks / hidden and change the count), bad descriptors come back as status codes, not exceptions"""
    import ctypes
    from gpflow_b200 import _lib
    lib = _lib.load()
    d = _lib.GpFlowFfjordDesc(abi_version=_lib.ABI_VERSION, dtype=_lib.F32, dim=10, n_hidden=2, n_blocks=3, n_steps=64, reverse=0)
    d.hidden[0], d.hidden[1] = 64, 64
    for i in range(3):
        d.t_end[i] = 1.0
    n = ctypes.c_int64(0)
    assert lib.gpflow_ffjord_param_count(ctypes.byref(d), ctypes.byref(n)) == 0 and n.value == 17784
    assert ctypes.sizeof(_lib.GpFlowFfjordDesc) == 4 * 4 + 7 * 4 + 3 * 4 + 8 * 8      # include/gpflow_b200.h: no padding
    d.n_blocks = 9
    assert lib.gpflow_ffjord_param_count(ctypes.byref(d), ctypes.byref(n)) == 1     # GPFLOW_ERR_BAD_DESC
    d.n_blocks, d.abi_version = 3, 99
    assert lib.gpflow_ffjord_param_count(ctypes.byref(d), ctypes.byref(n)) == 1
    d.abi_version = _lib.ABI_VERSION
    assert lib.gpflow_ffjord_integrate(ctypes.byref(d), ctypes.c_int64(4), None, None, None, None, None) == 3   # NULL pointer
    assert lib.gpflow_ffjord_integrate(ctypes.byref(d), ctypes.c_int64(0), None, None, None, None, None) == 0   # empty batch


# ---------------------------------------------------------------- GPU: the CUDA kernel
def _need_cuda():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")


def _map(g, spec, n_steps=None):
    import gpflow_b200 as G
    return G.FFJORDMap(spec.d, spec.hidden, spec.n_blocks, torch.from_numpy(g["params"]), spec.t_end,
                       n_steps or spec.n_steps).cuda()


@pytest.mark.gpu
@pytest.mark.parametrize("prec,tol", [("f32", 1e-5), ("f64", 1e-10)])
def test_cuda_matches_reference_golden(prec, tol):
    _need_cuda()
    g, spec = _golden(prec)
    m = _map(g, spec)
    x = torch.from_numpy(g["x"]).cuda()
    z, ld = m(x)
    assert _rel(z.cpu(), torch.from_numpy(g["z"])) < tol
    assert _rel(ld.cpu(), torch.from_numpy(g["logdet"])) < tol
    assert _rel(m.map_only(x).cpu(), torch.from_numpy(g["z"])) < tol
    xb = m.backward(torch.from_numpy(g["z"]).cuda())
    assert _rel(xb.cpu(), torch.from_numpy(g["xb"])) < tol


@pytest.mark.gpu
@pytest.mark.parametrize("B", [1, 31, 257, 4099])
def test_cuda_matches_fp64_oracle_ragged(B):
    _need_cuda()
    g, spec = _golden("f32")
    m = _map(g, spec)
    x = torch.randn(B, spec.d, generator=torch.Generator().manual_seed(B))
    z, ld = m(x.cuda())
    zo, ldo = FO.integrate(torch.from_numpy(g["params"]).double(), x.double(), spec, want_logdet=True)
    assert _rel(z.cpu().double(), zo) < 1e-5
    assert _rel(ld.cpu().double(), ldo) < 1e-5
    xb = m.backward(x.cuda())
    xo, _ = FO.integrate(torch.from_numpy(g["params"]).double(), x.double(), spec, reverse=True)
    assert _rel(xb.cpu().double(), xo) < 1e-5


@pytest.mark.gpu
def test_cuda_full_tiles_two_samples_per_thread():
    """a batch that fills every SM switches the fp32 kernel to two samples per thread: ragged last tile, fp64 oracle"""
    _need_cuda()
    g, spec = _golden("f32")
    B = 148 * 256 + 77
    spec.n_steps = 2
    m = _map(g, spec, n_steps=2)
    x = torch.randn(B, spec.d, generator=torch.Generator().manual_seed(5))
    fp = torch.from_numpy(g["params"]).double()
    z, ld = m(x.cuda())
    zo, ldo = FO.integrate(fp, x.double(), spec, want_logdet=True)
    assert _rel(z.cpu().double(), zo) < 1e-5
    assert _rel(ld.cpu().double(), ldo) < 1e-5
    xb = m.backward(x.cuda())
    xo, _ = FO.integrate(fp, x.double(), spec, reverse=True)
    assert _rel(xb.cpu().double(), xo) < 1e-5
    assert float((xb.cpu().double() - xo).abs().max()) < 1e-4


SHAPES = [  # d, hidden, blocks, steps
    (2, [], 1, 4), (2, [5], 2, 3), (3, [9, 17], 1, 5), (5, [8, 8, 8], 3, 2), (10, [64, 64], 1, 3), (17, [33], 2, 2),
    (7, [128, 40], 1, 2), (4, [16, 16, 16, 16], 1, 2)]


@pytest.mark.gpu
@pytest.mark.parametrize("d,hidden,blocks,steps", SHAPES)
@pytest.mark.parametrize("dtype,tol", [(torch.float32, 1e-5), (torch.float64, 1e-10)])
def test_cuda_shape_sweep_vs_oracle(d, hidden, blocks, steps, dtype, tol):
    _need_cuda()
    import gpflow_b200 as G
    spec = FO.FfjordSpec(d=d, hidden=hidden, n_blocks=blocks, n_steps=steps, t_end=[1.0, 0.7, 1.3][:blocks])
    flat = _rand_params(spec, 5 + d, dtype)
    x = torch.randn(77, d, generator=torch.Generator().manual_seed(d), dtype=torch.float64).to(dtype)
    m = G.FFJORDMap(d, hidden, blocks, flat, spec.t_end, steps).cuda()
    z, ld = m(x.cuda())
    zo, ldo = FO.integrate(flat.double(), x.double(), spec, want_logdet=True)
    assert _rel(z.cpu().double(), zo) < tol
    assert float((ld.cpu().double() - ldo).abs().max()) < tol * max(1.0, float(ldo.abs().max()))
    xb = m.backward(x.cuda())
    xo, _ = FO.integrate(flat.double(), x.double(), spec, reverse=True)
    assert _rel(xb.cpu().double(), xo) < tol


@pytest.mark.gpu
def test_gaussian_noise_dataset_round_trip_and_convergence():
    """size-independent properties at the reference's data-set size (n_samples = 3e5, latent_params.py:190):
    forward(backward(gauss)) == gauss up to the integrator error; 64 steps vs 128 steps agree below dopri5's tolerance."""
    _need_cuda()
    import gpflow_b200 as G
    g, spec = _golden("f32")
    m = _map(g, spec, n_steps=64)
    gen = torch.Generator(device="cuda").manual_seed(3)
    x_data, gauss = G.gaussian_noise_dataset(m, 300000, train=True, generator=gen)
    assert x_data.shape == (240000, spec.d) and x_data.is_cuda and bool(torch.isfinite(x_data).all())
    back = m.map_only(x_data)
    assert float((back - gauss).abs().max()) < 5e-4 and _rel(back, gauss) < 2e-5
    x128 = m.backward(gauss[:20000], n_steps=128)
    assert _rel(x_data[:20000], x128) < 2e-5
    x_val, _ = G.gaussian_noise_dataset(m, 300000, train=False, generator=gen)
    assert x_val.shape == (60000, spec.d)


@pytest.mark.gpu
def test_ffjord_errors():
    _need_cuda()
    import gpflow_b200 as G
    g, spec = _golden("f32")
    m = _map(g, spec)
    with pytest.raises(RuntimeError):
        m(torch.zeros(4, spec.d))                 # CPU tensor: no CPU path
    with pytest.raises(ValueError):
        m(torch.zeros(4, spec.d + 1).cuda())
    with pytest.raises(ValueError):
        G.FFJORDMap(spec.d, spec.hidden, spec.n_blocks, torch.zeros(5))
    z, ld = m(torch.zeros(0, spec.d).cuda())
    assert z.shape == (0, spec.d) and ld.shape == (0,)


@pytest.mark.gpu
def test_c3_pipeline_dataset_to_train_step():
    """C3 end to end on the device: Gaussian points -> FFJORDMap.backward (data set) -> FFJORDMap.map_only (x_map) ->
    fused GP-flow train step at the dsprites-latent shape (d=10, nb=9, H=[50,50,50], N=15), every stage against its oracle."""
    _need_cuda()
    import gpflow_b200 as G
    import gpflow_oracle as O
    g, fspec = _golden("f32")
    m = _map(g, fspec, n_steps=32)
    fspec.n_steps = 32
    gauss = torch.randn(192, fspec.d, generator=torch.Generator().manual_seed(21))
    x_data = m.backward(gauss.cuda())
    x_map = m.map_only(x_data)
    fp = torch.from_numpy(g["params"]).double()
    xo, _ = FO.integrate(fp, gauss.double(), fspec, reverse=True)
    assert _rel(x_data.cpu().double(), xo) < 1e-5
    zo, _ = FO.integrate(fp, x_data.cpu().double(), fspec)
    assert _rel(x_map.cpu().double(), zo) < 1e-5
    assert _rel(x_map.cpu(), gauss) < 1e-3            # round trip up to the 32-step integrator error
    spec = O.FlowSpec(d=10, nb=9, hidden=[50, 50, 50], n_steps=15)
    cfg = G.FlowConfig(d=10, nb=9, hidden=(50, 50, 50), n_steps=15)
    params = O.init_params(spec, 0.05, 0, torch.float32)
    res = G.fused_train_step(cfg, O.flatten_params(params).cuda(), x_data, x_map)
    ref = O.train_step([p.double() for p in params], x_data.cpu().double(), x_map.cpu().double(), spec)
    assert abs(float(res["scalars"][0]) - float(ref["ot"])) < 1e-5 * abs(float(ref["ot"]))
    assert _rel(res["grads"].cpu().double(), ref["grads"]) < 1e-5


@pytest.mark.gpu
def test_network_mirror_with_ffjord_map():
    """`GPFlowNetwork` with the FFJORD mirror as `self.map` (network.py:27,76-78,191-253): the fused step maps the batch
    itself (points only, no trace), `ot_map` returns map + flow log-determinants, and `ot_map(reverse=True)` -- which needs
    `map.backward` -- inverts it up to the integrators' round-trip errors."""
    _need_cuda()
    import gpflow_b200 as G
    import gpflow_oracle as O
    g, fspec = _golden("f32")
    m = _map(g, fspec, n_steps=32)
    spec = O.FlowSpec(d=10, nb=9, hidden=[50, 50, 50], n_steps=15)
    params = O.init_params(spec, 0.05, 0, torch.float32)
    flow = G.get_gp_flow(10, {"nb_neurons": [50, 50, 50], "nb_div_blocks": 9}, 15).cuda()
    with torch.no_grad():
        for p, q in zip(flow.parameters(), params):
            p.copy_(q.reshape(p.shape))
    net = G.GPFlowNetwork(flow, map_module=m)
    x = m.backward(torch.randn(96, 10, generator=torch.Generator().manual_seed(4)).cuda())
    loss, _ = net.fused_training_step(x)
    grads = torch.cat([p.grad.reshape(-1) for p in flow.parameters()]).clone()
    cfg = G.FlowConfig(d=10, nb=9, hidden=(50, 50, 50), n_steps=15)
    res = G.fused_train_step(cfg, O.flatten_params(params).cuda(), x, m.map_only(x))
    assert abs(float(loss) - float(res["scalars"][0])) < 1e-6 * abs(float(loss))
    assert _rel(grads.cpu(), res["grads"].cpu()) < 1e-6
    y, ld = net.ot_map(x, return_log_det=True)
    _, ld_map = m(x)
    assert y.shape == x.shape and ld.shape == (96,) and bool(torch.isfinite(ld).all())
    y2, ld_gp = net.gp_flow(m.map_only(x))
    assert _rel(y.cpu(), y2.cpu()) < 1e-6
    assert float((ld - (ld_map + ld_gp)).abs().max()) < 1e-4 * max(1.0, float(ld.abs().max()))
    xb = net.ot_map(y, reverse=True)
    assert _rel(xb.cpu(), x.cpu()) < 2e-3
