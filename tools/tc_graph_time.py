"""C3 shape on the tensor-core path at small batches: eager call against a captured CUDA graph of the same call."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))
import gpflow_oracle as O
import gpflow_b200 as G
spec = O.FlowSpec(d=10, nb=9, hidden=[50, 50, 50], n_steps=15)
flat = O.flatten_params(O.init_params(spec, 0.05, 0, torch.float32)).cuda()
cfg = G.FlowConfig(d=10, nb=9, hidden=(50, 50, 50), n_steps=15)
for B in (1024, 3072, 8192, 16384):
    xm = torch.randn(B, 10, device="cuda"); xd = 0.9 * xm.roll(1, 1) + 0.1
    out = torch.zeros(cfg.n_params + 3, device="cuda")
    kw = dict(force_tc=True, out=out)
    def timed(fn):
        ts = []
        for _ in range(5):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record(); e1.synchronize(); ts.append(e0.elapsed_time(e1))
        return sorted(ts)[2]
    eager = timed(lambda: G.fused_train_step(cfg, flat, xd, xm, **kw))
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        G.fused_train_step(cfg, flat, xd, xm, **kw)
    graph = timed(g.replay)
    print("B=%5d  eager %.2f ms   graph replay %.2f ms" % (B, eager, graph), flush=True)
