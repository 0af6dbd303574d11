# C3 shape: generic kernel against the tensor-core path around the dispatch threshold (gpflow_api.cu: tc_min_batch)
python - <<'PY'
import os, sys, torch
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), "oracle"))
import gpflow_oracle as O
import gpflow_b200 as G
spec = O.FlowSpec(d=10, nb=9, hidden=[50, 50, 50], n_steps=15)
flat = O.flatten_params(O.init_params(spec, 0.05, 0, torch.float32)).cuda()
cfg = G.FlowConfig(d=10, nb=9, hidden=(50, 50, 50), n_steps=15)
for B in (1024, 2048, 3072, 4096, 6144, 8192):
    xm = torch.randn(B, 10, device="cuda"); xd = 0.9 * xm.roll(1, 1) + 0.1
    out = []
    for kw in (dict(force_generic=True), dict(force_tc=True)):
        ts = []
        for i in range(5):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); G.fused_train_step(cfg, flat, xd, xm, **kw); e1.record(); e1.synchronize()
            ts.append(e0.elapsed_time(e1))
        out.append(sorted(ts)[2])
    print("B=%5d generic %.2f ms  tc %.2f ms" % (B, out[0], out[1]), flush=True)
PY
