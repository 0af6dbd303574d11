"""Stage-by-stage check of the chunked tensor-core path (force_tc) against the fp64 oracle and the generic kernel.
Run on a B200:  python tools/tc_debug.py [d nb h1,h2,.. B n_steps std]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))
import gpflow_oracle as O  # noqa: E402
import gpflow_b200 as G  # noqa: E402
from gpflow_b200 import ops_aux  # noqa: E402


def rel(a, b):
    a, b = a.double().cpu(), b.double().cpu()
    return float((a - b).norm() / max(float(b.norm()), 1e-300))


def run(d, nb, hidden, B, N, std, bench_cfg=False):
    print("=== d=%d nb=%d H=%s B=%d N=%d std=%g%s" % (d, nb, hidden, B, N, std, " (bench.py inputs)" if bench_cfg else ""), flush=True)
    spec = O.FlowSpec(d=d, nb=nb, hidden=list(hidden), n_steps=N)
    gen = torch.Generator().manual_seed(100 + d * nb)
    if bench_cfg:
        from bench import synth_batch
        params = [p.double() for p in O.init_params(spec, std, 0, torch.float32)]
        x_data, x_map = [t.double() for t in synth_batch(B, d, 1234, "cpu")]
    else:
        params = [p.float().double() for p in O.init_params(spec, std=std, seed=11 + d, dtype=torch.float64)]
        x_map = torch.randn(B, d, generator=gen, dtype=torch.float64).clamp(-2.5, 2.5).float().double()
        x_data = (1.1 * x_map.flip(-1) + 0.1).float().double()
    cfg = G.FlowConfig(d=d, nb=nb, hidden=tuple(hidden), n_steps=N)
    flat = O.flatten_params(params).float().cuda()
    # 1. one velocity evaluation
    xc = torch.erf(x_map / 2 ** 0.5)
    flow = G.get_gp_flow(d, {"nb_neurons": list(hidden), "nb_div_blocks": nb}, N).cuda()
    with torch.no_grad():
        off = 0
        for p in flow.parameters():
            p.copy_(flat[off:off + p.numel()].view_as(p)); off += p.numel()
    v_or = O.velocity(params, xc, torch.tensor(0.3, dtype=torch.float64), nb, O.eps_of(torch.float32))
    v_tc = ops_aux.velocity(flow.v, xc.float().cuda(), 0.3, force_tc=True)
    v_ge = ops_aux.velocity(flow.v, xc.float().cuda(), 0.3)
    print("velocity: tc vs oracle %.3e   generic vs oracle %.3e" % (rel(v_tc, v_or), rel(v_ge, v_or)), flush=True)
    # 2. forward op
    r = O.train_step(params, x_data, x_map, spec, with_lagrangian=True, need_grad_x=True)
    f_tc = G.flow_forward(cfg, flat, x_map.float().cuda(), wrap=True, lagrangian=True, want_logdet=True, want_traj=True, force_tc=True)
    f_ge = G.flow_forward(cfg, flat, x_map.float().cuda(), wrap=True, lagrangian=True, want_logdet=True, want_traj=True, force_generic=True)
    print("forward y: tc %.3e generic %.3e | traj tc-vs-generic %.3e | lagr tc %.6e generic %.6e oracle %.6e" % (
        rel(f_tc["y"], r["y"]), rel(f_ge["y"], r["y"]), rel(f_tc["traj"], f_ge["traj"]), float(f_tc["scalars"][2]),
        float(f_ge["scalars"][2]), float(r["lagrangian"])), flush=True)
    # 3. train step
    t_tc = G.fused_train_step(cfg, flat, x_data.float().cuda(), x_map.float().cuda(), want_y=True, want_grad_x=True, force_tc=True)
    t_ge = G.fused_train_step(cfg, flat, x_data.float().cuda(), x_map.float().cuda(), want_y=True, want_grad_x=True, force_generic=True)
    print("train: ot tc %.8e generic %.8e oracle %.8e" % (float(t_tc["scalars"][0]), float(t_ge["scalars"][0]), float(r["ot"])))
    print("       grads tc %.3e generic %.3e | grad_x tc %.3e generic %.3e" % (
        rel(t_tc["grads"], r["grads"]), rel(t_ge["grads"], r["grads"]), rel(t_tc["grad_x"], r["grad_x"]),
        rel(t_ge["grad_x"], r["grad_x"])), flush=True)
    s = spec.layer_sizes()
    off = 0
    for i in range(len(s) - 1):
        nW, nb_ = s[i] * s[i + 1], s[i + 1]
        print("       layer %d: W %.3e (|g| %.3e)  b %.3e (|g| %.3e)" % (
            i, rel(t_tc["grads"][off:off + nW], r["grads"][off:off + nW]), float(r["grads"][off:off + nW].norm()),
            rel(t_tc["grads"][off + nW:off + nW + nb_], r["grads"][off + nW:off + nW + nb_]),
            float(r["grads"][off + nW:off + nW + nb_].norm())), flush=True)
        off += nW + nb_
    # 4. backward op
    gy = torch.randn(B, d, generator=gen, dtype=torch.float64).float()
    b_tc = G.flow_backward(cfg, flat, x_map.float().cuda(), gy.cuda(), wrap=True, force_tc=True)
    b_ge = G.flow_backward(cfg, flat, x_map.float().cuda(), gy.cuda(), wrap=True, force_generic=True)
    print("backward op: grads tc-vs-generic %.3e  grad_x %.3e" % (rel(b_tc[0], b_ge[0]), rel(b_tc[1], b_ge[1])), flush=True)


if __name__ == "__main__":
    if len(sys.argv) > 1:
        d, nb = int(sys.argv[1]), int(sys.argv[2])
        hidden = tuple(int(h) for h in sys.argv[3].split(","))
        run(d, nb, hidden, int(sys.argv[4]), int(sys.argv[5]), float(sys.argv[6]))
    else:
        run(5, 3, (64, 64), 300, 3, 0.05)
        run(6, 6, (40,), 130, 3, 0.05)
        run(10, 9, (50, 50, 50), 200, 3, 0.03)
        run(17, 16, (100, 70), 260, 2, 0.02)
        os.environ["GPFLOW_TC_SLAB"] = "256"       # three slabs: accumulation across slabs
        run(10, 9, (50, 50, 50), 700, 3, 0.03)
        del os.environ["GPFLOW_TC_SLAB"]
        run(10, 9, (50, 50, 50), 64, 15, 0.05, bench_cfg=True)
        run(64, 63, (256, 256), 8, 15, 0.02, bench_cfg=True)
