"""Standalone `VectorField.eval / .jacobian / __call__` on the GPU (`gpflow_potentials`, csrc/gpflow_potentials.cu):
the reference's own test of this class ported (tests/test_vector_field.py:15-25: the Jacobian equals the autograd gradient
of eval, dim 4 -> 1, [5, 5, 5], 50 points, 5e-6), plus the oracle's potentials (A.1) for the shapes of the path, both
precisions, the time column, a 1-D input and the size limit."""
import pytest
import torch

import gpflow_oracle as O
from golden_util import rel

pytestmark = pytest.mark.gpu


def _torch_mlp(vf, x):
    """The same network in plain torch on the CPU (fp64), differentiable: the check, not the product."""
    h = x
    layers = [(d.weight.detach().double().cpu(), d.bias.detach().double().cpu()) for d in vf.denses]
    for i, (W, b) in enumerate(layers):
        h = h @ W.T + b
        if i != len(layers) - 1:
            h = torch.tanh(h)
    return h


def test_reference_test_grad_ported():
    from gpflow_b200.vector_field import VectorField
    torch.manual_seed(0)
    dim = 4
    v = VectorField(dim, 1, nb_neurons=[5, 5, 5]).cuda()
    with torch.no_grad():
        for d in v.denses:                       # the reference initialises N(0, 1e-3): use O(1) weights so that the test bites
            d.weight.normal_(0.0, 0.7)
            d.bias.normal_(0.0, 0.3)
    x = torch.rand(50, dim)
    xg = x.double().requires_grad_(True)
    v_x = _torch_mlp(v, xg)
    grad1 = torch.autograd.grad(v_x, xg, torch.ones_like(v_x))[0]
    grad2 = v.jacobian(x.cuda()).squeeze(1).cpu().double()
    assert float(torch.norm(grad1 - grad2)) < 5e-6                      # the reference's assertion
    assert rel(v.eval(x.cuda()).cpu(), v_x.detach()) < 1e-6
    assert torch.equal(v(x.cuda()), v.eval(x.cuda()))                   # __call__ is eval (vector_field.py:72-73)
    j1 = v.jacobian(x[3].cuda())                                        # 1-D input (vector_field.py:103)
    assert j1.shape == (1, dim) and rel(j1.cpu(), grad1[3:4]) < 1e-5


@pytest.mark.parametrize("d,nb,hidden,dtype,tol", [
    (2, 1, [15, 15], torch.float32, 1e-5), (2, 1, [15, 15], torch.float64, 1e-10),
    (10, 9, [50, 50, 50], torch.float32, 1e-5), (5, 3, [5, 5, 5], torch.float64, 1e-10),
    (64, 63, [256, 256], torch.float32, 1e-5),
])
def test_potentials_match_the_oracle(d, nb, hidden, dtype, tol):
    """The velocity field's network: dim_in = d + 1 ([x; t]), dim_out = d nb (velocity_field.py:88-95)."""
    import gpflow_b200 as G
    vf = G.VelocityField(d, list(hidden), nb_div_blocks=nb, dtype=dtype).cuda()
    net = vf.v
    spec = O.FlowSpec(d=d, nb=nb, hidden=list(hidden), n_steps=1)
    params = O.init_params(spec, 0.5 if d < 64 else 0.05, 3, dtype)
    with torch.no_grad():
        for p, q in zip(net.parameters(), params):
            p.copy_(q)
    B, t = 37, 0.4
    x = (2 * torch.rand(B, d, dtype=torch.float64) - 1).to(dtype)
    xt = torch.cat([x, torch.full((B, 1), t, dtype=dtype)], dim=1).cuda()
    f, J = net.eval(xt), net.jacobian(xt)
    assert f.shape == (B, d * nb) and J.shape == (B, d * nb, d + 1)
    f_ref, J_ref = O.potentials([p.double() for p in params], x.double(), torch.tensor(t, dtype=torch.float64))
    assert rel(f.cpu(), f_ref) < tol, rel(f.cpu(), f_ref)
    assert rel(J[..., :d].cpu(), J_ref) < tol, rel(J[..., :d].cpu(), J_ref)
    # the time column against a central difference of the oracle in fp64
    h = 1e-6
    fp, _ = O.potentials([p.double() for p in params], x.double(), torch.tensor(t + h, dtype=torch.float64))
    fm, _ = O.potentials([p.double() for p in params], x.double(), torch.tensor(t - h, dtype=torch.float64))
    assert rel(J[..., d].cpu(), (fp - fm) / (2 * h)) < max(tol, 1e-7) * 10


def test_potentials_size_limit_and_errors():
    import ctypes
    from gpflow_b200 import _lib
    from gpflow_b200.vector_field import VectorField
    lib = _lib.load()
    big = VectorField(65, 64 * 63, nb_neurons=[256, 256]).cuda().double()
    x = torch.zeros(2, 65, dtype=torch.float64, device="cuda")
    with pytest.raises(RuntimeError, match="not supported"):          # 2 x 256 x 66 doubles > 200 KB of shared memory
        big.eval(x)
    w = (ctypes.c_int32 * 3)(4, 5, 1)
    assert lib.gpflow_potentials(0, 2, w, 4, None, None, None, None, None) == 3          # NULL parameters
    assert lib.gpflow_potentials(7, 2, w, 4, ctypes.c_void_p(8), ctypes.c_void_p(8), None, None, None) == 1
    with pytest.raises(ValueError):
        VectorField(4, 1, nb_neurons=[5]).cuda().eval(torch.zeros(3, 5, device="cuda"))
