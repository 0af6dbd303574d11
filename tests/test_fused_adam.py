"""Fused Adam(+L2) + StepLR on the flat parameter buffer against torch.optim.Adam / StepLR,
the optimiser the reference builds in `Network.configure_optimizers` (src/gp_flows/network.py:307-320)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _need_cuda():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 2e-6), (torch.float64, 1e-13)])
def test_adam_kernel_matches_torch_adam(dtype, tol):
    _need_cuda()
    import gpflow_b200 as G
    torch.manual_seed(3)
    shapes = [(15, 3), (15,), (15, 15), (15,), (2, 15), (2,)]
    ref_p = [torch.nn.Parameter(torch.randn(s, device="cuda", dtype=dtype)) for s in shapes]
    our_p = [torch.nn.Parameter(p.detach().clone()) for p in ref_p]
    ref = torch.optim.Adam(ref_p, lr=2e-3, betas=(0.9, 0.999), weight_decay=1e-4)
    sched = torch.optim.lr_scheduler.StepLR(ref, step_size=3, gamma=0.5)
    ours = G.FlatAdam(our_p, lr=2e-3, betas=(0.9, 0.999), weight_decay=1e-4, step_size=3, gamma=0.5)
    assert ours.intact()
    for epoch in range(7):
        for it in range(4):
            grads = [torch.randn(s, device="cuda", dtype=dtype) * (0.1 + it) for s in shapes]
            for p, q, g in zip(ref_p, our_p, grads):
                p.grad = g.clone()
                q.grad.copy_(g)          # views of the flat gradient buffer
            ref.step()
            ours.step()
        sched.step()
        ours.scheduler_step()
        assert abs(ours.lr - ref.param_groups[0]["lr"]) < 1e-15
    for p, q in zip(ref_p, our_p):
        err = (p.detach() - q.detach()).abs().max().item() / max(p.detach().abs().max().item(), 1e-30)
        assert err < tol, err
    st = ref.state[ref_p[2]]
    off = 15 * 3 + 15
    m = ours.exp_avg[off:off + 225].view(15, 15)
    assert (st["exp_avg"] - m).abs().max().item() < tol * max(1.0, st["exp_avg"].abs().max().item())


def test_fused_training_step_in_place_matches_copy_path():
    """GPFlowNetwork.fused_training_step + FlatAdam (in-place flat buffers) == the same steps with torch Adam."""
    _need_cuda()
    import copy
    import gpflow_b200 as G
    torch.manual_seed(0)
    flow_a = G.get_gp_flow(2, {"nb_neurons": [15, 15]}, 15).cuda()
    with torch.no_grad():
        for p in flow_a.parameters():
            p.normal_(0.0, 0.3)
    flow_b = copy.deepcopy(flow_a)
    net_a, net_b = G.GPFlowNetwork(flow_a), G.GPFlowNetwork(flow_b)
    opt_a = torch.optim.Adam(flow_a.parameters(), lr=2e-3, betas=(0.9, 0.999), weight_decay=1e-4)
    opt_b = net_b.configure_fused_optimizer(lr=2e-3, weight_decay=1e-4)
    assert opt_b.intact()
    g = torch.Generator(device="cuda").manual_seed(5)
    for k in range(6):
        x_map = torch.randn(2048, 2, generator=g, device="cuda")
        x = 1.3 * x_map + 0.2
        la, _ = net_a.fused_training_step(x, x_map)
        opt_a.step()
        lb, _ = net_b.fused_training_step(x, x_map)
        opt_b.step()
        assert abs(float(la) - float(lb)) <= 1e-5 * abs(float(la)), (k, float(la), float(lb))
    pa = torch.cat([p.detach().reshape(-1) for p in flow_a.parameters()])
    pb = torch.cat([p.detach().reshape(-1) for p in flow_b.parameters()])
    assert opt_b.intact() and pb.data_ptr() != opt_b.flat.data_ptr()
    assert (pa - pb).abs().max().item() < 2e-5 * pa.abs().max().item()
    assert (pb - opt_b.flat).abs().max().item() == 0.0
    sd = flow_b.state_dict()
    assert "v.v.denses.0.weight" in sd          # reference checkpoint keys survive the re-homing
