"""Timing of the fused train step through the chunked tensor-core path vs the generic kernel (CUDA events).
    python tools/tc_time.py c5|c3 log2B [steps] [tc|generic]"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))
import gpflow_oracle as O  # noqa: E402
import gpflow_b200 as G  # noqa: E402
from bench import WORKLOADS, synth_batch  # noqa: E402

name = {"c5": "c5_d64_nb63", "c3": "c3_d10_nb9"}[sys.argv[1]]
B = 1 << int(sys.argv[2])
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
which = sys.argv[4] if len(sys.argv) > 4 else "tc"
wl = WORKLOADS[name]
spec = O.FlowSpec(d=wl["d"], nb=wl["nb"], hidden=list(wl["hidden"]), n_steps=wl["n_steps"])
flat = O.flatten_params(O.init_params(spec, wl["w_std"], 0, torch.float32)).cuda()
cfg = G.FlowConfig(d=wl["d"], nb=wl["nb"], hidden=tuple(wl["hidden"]), n_steps=wl["n_steps"])
x_data, x_map = synth_batch(B, wl["d"], 1234, "cuda")
kw = dict(force_tc=True) if which == "tc" else dict(force_generic=True)
out = None
for i in range(1 + steps):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    res = G.fused_train_step(cfg, flat, x_data, x_map, **kw)
    e1.record()
    t_host = time.perf_counter() - t0
    e1.synchronize()
    ms = e0.elapsed_time(e1)
    print("%s %s B=%d step %d: %.2f ms (host enqueue %.2f ms) -> %.1f samples/s   ot %.6f |g| %.5e" % (
        name, which, B, i, ms, t_host * 1e3, B / ms * 1e3, float(res["scalars"][0]), float(res["grads"].norm())), flush=True)
print("max memory allocated: %.1f MB" % (torch.cuda.max_memory_allocated() / 2 ** 20))
