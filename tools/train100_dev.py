"""Deviation of the CUDA path from the reference's 100-step training trajectories (tests/golden/train100_*): prints the
largest per-step relative OT-cost error and the relative parameter error after 100 Adam steps."""
import os, sys
import numpy as np, torch
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import gpflow_b200 as G
GD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")
for name in ("train100_c1_eight_gaussians", "train100_c2_moons_euler", "train100_c3_latent_synthetic"):
    z = np.load(os.path.join(GD, name + "_f32.npz")); g = {k: z[k] for k in z.files}
    euler = None if float(g["euler_coef"]) < 0 else float(g["euler_coef"])
    d, nb, hidden = int(g["d"]), int(g["nb"]), [int(h) for h in g["hidden"]]
    flow = G.get_gp_flow(d, {"nb_neurons": hidden, "nb_div_blocks": nb}, 15,
                         euler_dict={"use_euler": euler is not None, "case": "penalization", "coef_penalization": euler}).cuda()
    off, flat0 = 0, torch.from_numpy(g["params0"])
    with torch.no_grad():
        for p in flow.parameters():
            p.copy_(flat0[off:off + p.numel()].view_as(p)); off += p.numel()
    net = G.GPFlowNetwork(flow, map_module=None)
    opt = torch.optim.Adam(flow.parameters(), lr=2e-3, betas=(0.9, 0.999), weight_decay=1e-4)
    worst, worst_reg = 0.0, 0.0
    for k in range(int(g["K"])):
        torch.manual_seed(5000 + k)
        loss, reg = net.fused_training_step(torch.from_numpy(g["x"][k]).cuda(), torch.from_numpy(g["x_map"][k]).cuda())
        worst = max(worst, abs(float(loss) - g["ot"][k]) / abs(g["ot"][k]))
        worst_reg = max(worst_reg, abs(float(reg) - g["reg"][k]) / abs(g["ot"][k]))
        opt.step()
    flat = torch.cat([p.detach().reshape(-1) for p in flow.parameters()]).cpu().double()
    ref = torch.from_numpy(g["paramsK"]).double()
    print(name, G.kernel_variant(flow.config(), euler_coef=euler), "max ot rel err %.2e, max |reg err|/ot %.2e, params rel err %.2e, |dparams|/|params| %.2e" % (
        worst, worst_reg, float((flat - ref).norm() / ref.norm()), float((ref - flat0.double()).norm() / ref.norm())))
