"""Run-to-run reproducibility of the tc train step: which part of [grads | scalars] differs between two identical calls."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))
import gpflow_oracle as O
import gpflow_b200 as G

for d, nb, hidden, B, N in [(5, 3, (64, 64), 300, 2), (6, 6, (40,), 130, 2), (10, 9, (50, 50, 50), 200, 2)]:
    spec = O.FlowSpec(d=d, nb=nb, hidden=list(hidden), n_steps=N)
    params = O.init_params(spec, 0.05, 3, torch.float32)
    x_map = torch.randn(B, d, generator=torch.Generator().manual_seed(1)).clamp(-2.5, 2.5)
    x_data = 1.1 * x_map.flip(-1) + 0.1
    cfg = G.FlowConfig(d=d, nb=nb, hidden=tuple(hidden), n_steps=N)
    flat = O.flatten_params(params).cuda()
    outs = []
    for it in range(4):
        r = G.fused_train_step(cfg, flat, x_data.cuda(), x_map.cuda(), want_y=True, want_grad_x=True, force_tc=True)
        outs.append((r["buf"].clone(), r["y"].clone(), r["grad_x"].clone()))
        torch.empty(1 << 26, device="cuda").normal_()      # dirty the allocator's free blocks
    s = spec.layer_sizes()
    print("shape", d, nb, hidden)
    for it in range(1, 4):
        print("  run %d: y diff %.3e grad_x diff %.3e scalars diff %.3e" % (
            it, float((outs[it][1] - outs[0][1]).abs().max()), float((outs[it][2] - outs[0][2]).abs().max()),
            float((outs[it][0][-3:] - outs[0][0][-3:]).abs().max())))
        off = 0
        for i in range(len(s) - 1):
            nW, nb_ = s[i] * s[i + 1], s[i + 1]
            dw = (outs[it][0][off:off + nW] - outs[0][0][off:off + nW]).abs()
            db = (outs[it][0][off + nW:off + nW + nb_] - outs[0][0][off + nW:off + nW + nb_]).abs()
            print("     layer %d: W max diff %.3e (%d entries differ)   b max diff %.3e (%d)" % (
                i, float(dw.max()), int((dw > 0).sum()), float(db.max()), int((db > 0).sum())))
            off += nW + nb_
