"""TEST INFRASTRUCTURE ONLY -- golden vectors of the frozen BNAF maps, produced by the UNMODIFIED reference modules
(needs /root/reference):

    python oracle/make_golden_bnaf.py

For pretrained_models/{eight_gaussians,moons}/bnaf/torch_nf_model.pt (flow_models/bnaf/bnaf.py) stores the flat
parameters ([_weight | _diag_weight | bias] per masked layer), 384 inputs (256 samples of the data set via
`inf_train_gen`, src/datasets/distributions_toy.py:12-85, plus 128 wide Gaussian points) and the module's own
`(y, log_det)` in fp32 -> tests/golden/bnaf_<data>_f32.npz.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")
REF = "/root/reference"


def main():
    sys.path.insert(0, HERE)
    import torch
    import ref_import
    ref_import.import_reference(fp64=False)
    from src.datasets.distributions_toy import inf_train_gen
    for data in ("eight_gaussians", "moons"):
        nf = torch.load(os.path.join(REF, "pretrained_models", data, "bnaf", "torch_nf_model.pt"), weights_only=False,
                        map_location="cpu")
        nf.eval()
        sd = nf.state_dict()
        idx = sorted({int(k.split(".")[2]) for k in sd if k.endswith("._weight")})
        flat = torch.cat([torch.cat([sd["flow.0.%d._weight" % i].reshape(-1), sd["flow.0.%d._diag_weight" % i].reshape(-1),
                                     sd["flow.0.%d.bias" % i].reshape(-1)]) for i in idx])
        rng = np.random.RandomState(7)
        x = np.concatenate([inf_train_gen(data, rng=rng, batch_size=256).astype(np.float32),
                            (2.5 * rng.randn(128, 2)).astype(np.float32)])
        with torch.no_grad():
            y, ld = nf(torch.from_numpy(x))
        first = sd["flow.0.%d._weight" % idx[0]]
        path = os.path.join(OUT, "bnaf_%s_f32.npz" % data)
        np.savez_compressed(path, params=flat.numpy(), x=x, y=y.numpy(), logdet=ld.numpy(), dim=int(first.shape[1]),
                            n_neurons=int(first.shape[0]), n_layers=len(idx))
        print("wrote", path, os.path.getsize(path), "bytes; params", flat.numel(), "y range", float(y.min()), float(y.max()))


if __name__ == "__main__":
    main()
