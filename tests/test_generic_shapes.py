"""Generic kernel, shape sweep against the fp64 oracle on identical seeded inputs (no golden file needed: the oracle is
pinned by tests/golden).  The shapes are chosen to reach every branch of the fp32 fast paths of gpflow_generic.cuh /
gpflow_generic_mma.cuh: ragged mma tiles (rows 1 + d not a multiple of 16, widths not multiples of 8), the 128-bit
last-layer passes on and off (last hidden width / parameter offset multiple of 4 or not), the on-chip (shared memory)
and the global scratch placement, nb = 1 ... d, one to three hidden layers, batches below / above the grid size.
Bars: 1e-5 (fp32) / 1e-10 (fp64) on y, OT cost, Lagrangian and parameter gradients for every shape; 2e-5 on d loss / d x_map,
which carries the erfinv conditioning factor of the wrapper (tests/test_sf2_parity.py: GX_TOL)."""
import pytest
import torch

import gpflow_oracle as O
from golden_util import rel

pytestmark = pytest.mark.gpu

SHAPES = [
    # d, nb, hidden, B, weight std (tuned so that the flow moves the points by ~25-35 % of their norm)
    (3, 2, (20, 12), 37, 0.08),          # vec4 (H_L = 12, offset 352), on-chip
    (4, 3, (8, 8, 8), 5, 0.08),          # three layers, vec4, tiny batch
    (6, 5, (16,), 64, 0.05),             # single hidden layer (no mma layer), vec4
    (5, 1, (7, 9), 33, 0.12),            # odd widths: scalar last layer, ragged mma tiles
    (17, 16, (24, 40), 19, 0.02),        # 18 rows = two m-tiles, nb = d - 1
    (12, 12, (33, 31), 9, 0.03),         # nb = d, odd widths
    (20, 4, (128, 128), 300, 0.012),     # does not fit on chip -> global scratch; vec4; several n-blocks of 64 columns
    (2, 1, (15, 15), 70, 0.12),          # the sf2 shape through the generic kernel
    (3, 2, (64, 64), 150, 0.04),         # aligned mma fast path with P % 4 == 2 (regression: split copies misaligned)
    (10, 9, (64, 64, 64), 40, 0.03),     # the zero-padded C3 shape: fast path, P = 14938
    (10, 9, (50, 50, 50), 40, 0.03),     # C3: runs zero-padded to [64,64,64] (fp32), gradients gathered back
    (6, 3, (50, 100), 33, 0.04),         # zero-padded to [64,128] (fp32)
]


@pytest.mark.parametrize("d,nb,hidden,B,std", SHAPES)
@pytest.mark.parametrize("dtype,tol", [(torch.float32, 1e-5), (torch.float64, 1e-10)])
def test_train_step_shape_sweep(d, nb, hidden, B, std, dtype, tol):
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    import gpflow_b200 as G
    spec = O.FlowSpec(d=d, nb=nb, hidden=list(hidden), n_steps=6)
    params = O.init_params(spec, std=std, seed=11 + d, dtype=torch.float64)
    gen = torch.Generator().manual_seed(100 + d * nb)
    x_map = torch.randn(B, d, generator=gen, dtype=torch.float64).clamp(-2.5, 2.5)
    x_data = 1.1 * x_map.flip(-1) + 0.1
    if dtype == torch.float32:       # identical inputs: round once, promote for the fp64 truth
        params = [p.float().double() for p in params]
        x_map, x_data = x_map.float().double(), x_data.float().double()
    r = O.train_step(params, x_data, x_map, spec, with_lagrangian=True, need_grad_x=True)
    cfg = G.FlowConfig(d=d, nb=nb, hidden=tuple(hidden), n_steps=6)
    flat = O.flatten_params(params).to(dtype).cuda()
    res = G.fused_train_step(cfg, flat, x_data.to(dtype).cuda(), x_map.to(dtype).cuda(), lagrangian=True, want_y=True,
                             want_grad_x=True, force_generic=True)
    sc = res["scalars"].cpu().double()
    assert rel(res["y"].cpu(), r["y"]) < tol, ("y", rel(res["y"].cpu(), r["y"]))
    assert abs(sc[0] - r["ot"]) < tol * abs(r["ot"]), ("ot", float(sc[0]), float(r["ot"]))
    assert abs(sc[2] - r["lagrangian"]) < tol * abs(r["lagrangian"]) + 1e-12
    assert rel(res["grads"].cpu(), r["grads"]) < tol, ("grads", rel(res["grads"].cpu(), r["grads"]))
    assert rel(res["grad_x"].cpu(), r["grad_x"]) < 2 * tol, ("grad_x", rel(res["grad_x"].cpu(), r["grad_x"]))
