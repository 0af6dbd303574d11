"""Calibration of bench.py's CPU reference arm (`kind: "port"`): the UNMODIFIED reference (imported from /root/reference
through oracle/ref_import.py) and the oracle port (oracle/gpflow_oracle.py) timed on identical batches on the same host
cores.  One step = loss_gp_flow + regularization_coef + backward (network.py:70-84,104-111), frozen map excluded.
Run in the build container:  python tools/calibrate_cpu_baseline.py   (prints a table for BASELINE.md section 2)"""
import os
import sys
import time
import types

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import gpflow_oracle as O  # noqa: E402
import ref_import  # noqa: E402
from bench import synth_batch  # noqa: E402


def main():
    rn = ref_import.import_reference()
    from src.gp_flows.network import Network
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    print("host: %d threads, torch %s" % (threads, torch.__version__))
    cases = [("C1/C4 shape d=2 H=[15,15]", 2, 1, [15, 15], 0.3, [1024, 65536]),
             ("C3 shape d=10 nb=9 H=[50,50,50]", 10, 9, [50, 50, 50], 0.05, [1024])]
    for name, d, nb, H, std, batches in cases:
        spec = O.FlowSpec(d=d, nb=nb, hidden=H, n_steps=15)
        ps = O.init_params(spec, std, 0, torch.float32)
        flow = rn.get_gp_flow(d, {"nb_neurons": list(H), "nb_div_blocks": nb}, 15)
        with torch.no_grad():
            for p, q in zip(flow.parameters(), ps):
                p.copy_(q)
        for B in batches:
            x_data, x_map = synth_batch(B, d, 1234, "cpu")
            ns = types.SimpleNamespace(flow=flow, eps=O.eps_of(torch.float32), sigma=1)

            def ref_step():
                y, _ = Network.gp_flow(ns, x_map)
                loss = ((x_data - y) ** 2).sum(-1).mean() + flow.regularization_coef()
                flow.zero_grad()
                loss.backward()
                return float(loss)

            def port_step():
                return float(O.train_step(ps, x_data, x_map, spec)["ot"])

            out = []
            for fn in (ref_step, port_step):
                fn()
                ts = []
                for _ in range(3):
                    t0 = time.perf_counter(); v = fn(); ts.append(time.perf_counter() - t0)
                ts.sort()
                out.append((ts[1], v))
            (tr, vr), (tp, vp) = out
            print("%-34s B=%6d  reference %8.1f ms (%8.0f samples/s)   port %8.1f ms (%8.0f samples/s)   port/reference speed %.2fx   "
                  "loss ref %.6f port %.6f" % (name, B, tr * 1e3, B / tr, tp * 1e3, B / tp, tr / tp, vr, vp), flush=True)


if __name__ == "__main__":
    main()
