"""developer aid: device time of the frozen FFJORD map (gpflow_ffjord_integrate) at the reference's data-set size
(n_samples = 3e5 -> 240000 training points, pretrained dsprites map: 3 blocks of 10 -> 64 -> 64 -> 10, 64 RK4 steps)."""
import os, sys, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpflow_b200 as G
z = np.load(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "ffjord_dsprites_f32.npz"))
hidden = [int(h) for h in z["hidden"]]
d, nb = int(z["d"]), int(z["n_blocks"])
m = G.FFJORDMap(d, hidden, nb, torch.from_numpy(z["params"]), [float(t) for t in z["t_end"]], 64).cuda()
n = [d] + hidden + [d]
macs = sum(n[i] * n[i + 1] for i in range(len(n) - 1))
tr = d * (sum(n[i] * n[i + 1] for i in range(1, len(n) - 2)) + n[1] + n[-2])
def timeit(fn, x, reps=5):
    for _ in range(2): fn(x)
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(x); e1.record(); e1.synchronize(); ts.append(e0.elapsed_time(e1))
    ts.sort(); return ts[len(ts) // 2]
for B in (2000, 240000, 1 << 20):
    x = torch.randn(B, d, device="cuda")
    evals = 4 * 64 * nb
    t = timeit(m.backward, x)
    print("backward  B=%7d  %.3f ms  %.3g samples/s  %.2f TFLOP/s (value only, %d MAC/eval)" % (B, t, B / t * 1e3, 2 * macs * evals * B / t * 1e-9, macs), flush=True)
    t = timeit(m.map_only, x)
    print("map_only  B=%7d  %.3f ms  %.3g samples/s" % (B, t, B / t * 1e3), flush=True)
    t = timeit(lambda v: m(v), x, reps=3)
    print("forward+logdet B=%7d  %.3f ms  %.3g samples/s  %.2f TFLOP/s (value + trace, %d MAC/eval)" % (B, t, B / t * 1e3, 2 * (macs + tr) * evals * B / t * 1e-9, macs + tr), flush=True)
