"""Helpers shared by the parity tests: load tests/golden/*.npz (outputs of the UNMODIFIED
reference, written by oracle/make_golden.py) and rebuild the inputs they were computed from."""
import glob
import os

import numpy as np
import torch

import gpflow_oracle as O

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def golden_names(tag):
    """single-step cases (oracle/make_golden.py); the train_* files are training trajectories (make_golden_train.py),
    the bnaf_* / ffjord_* files the frozen-map vectors (make_golden_bnaf.py, make_golden_ffjord.py), compose_* the
    nb_blocks != 1 vectors (make_golden_compose.py), measured_* the bench-configuration vectors (make_golden_measured.py)"""
    names = sorted(os.path.basename(p)[: -len("_%s.npz" % tag)] for p in glob.glob(os.path.join(GOLDEN_DIR, "*_%s.npz" % tag)))
    return [n for n in names if not n.startswith(("train_", "train100_", "bnaf_", "ffjord_", "compose_", "measured_", "ot_costs_"))]


def load_golden(name, tag):
    z = np.load(os.path.join(GOLDEN_DIR, "%s_%s.npz" % (name, tag)), allow_pickle=False)
    g = {k: z[k] for k in z.files}
    dtype = torch.float64 if tag == "f64" else torch.float32
    spec = O.FlowSpec(d=int(g["d"]), nb=int(g["nb"]), hidden=[int(h) for h in g["hidden"]], n_steps=int(g["N"]),
                      n_steps_eval=int(g["n_eval"]))
    if str(g["w_kind"]) == "formula":
        params = O.formula_params(spec, float(g["w_scale"]), dtype)
    else:
        flat = torch.from_numpy(g["params"]).to(dtype)
        params, off = [], 0
        s = spec.layer_sizes()
        for i in range(len(s) - 1):
            params.append(flat[off: off + s[i] * s[i + 1]].reshape(s[i + 1], s[i]).clone()); off += s[i] * s[i + 1]
            params.append(flat[off: off + s[i + 1]].clone()); off += s[i + 1]
    euler = float(g["euler_coef"])
    return dict(g=g, spec=spec, params=params, dtype=dtype, euler=None if euler < 0 else euler,
                x_map=torch.from_numpy(g["x_map"]).to(dtype), x_data=torch.from_numpy(g["x_data"]).to(dtype),
                zw=torch.from_numpy(g["zw"]).to(dtype))


def rel(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    den = np.linalg.norm(b.ravel())
    return float(np.linalg.norm((a - b).ravel()) / (den if den > 0 else 1.0))
