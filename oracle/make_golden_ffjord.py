"""TEST INFRASTRUCTURE ONLY -- golden vectors of the frozen FFJORD map, produced by the UNMODIFIED reference modules
(needs /root/reference):

    python oracle/make_golden_ffjord.py

pretrained_models/dsprites/ffjord/torch_nf_model.pt (flow_models/ffjord/ffjord.py: 3 CNF blocks, ODEnet 10 -> 64 -> 64
-> 10 concatsquash/tanh) is loaded and put in eval mode (brute-force divergence).  Two kinds of vectors are stored in
tests/golden/ffjord_dsprites_{f32,f64}.npz:

  * `vf_*`: the reference's own `ODEfunc.forward(t, (y, logp))` = (f(t,y), -div f) of every block at a few times
    (pure reference arithmetic: no integrator involved);
  * `z, logdet = nf(x)` and `xb = nf.backward(z)`: the reference's FFJORD / SequentialFlow / CNF / ODEfunc code path with
    the one thing that is missing in this image -- torchdiffeq 0.2.3's `odeint_adjoint` (cnf.py:4) -- replaced by a
    fixed-grid classical RK4 with `N_STEPS` steps per block (the `rk4_fixed` below, injected as
    `flow_models.ffjord.layers.cnf.odeint`).  The reference's adaptive dopri5 (atol = rtol = 1e-5) cannot be run here:
    parity with it is unpinned; `err_halving` records |z(N) - z(2N)| as the integrator's own error.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")
REF = "/root/reference"
N_STEPS = 16


def rk4_fixed_factory(n_steps):
    import torch

    def rk4_fixed(func, y0, t, atol=None, rtol=None, method=None, options=None):
        """torchdiffeq.odeint's calling convention for a tuple state and len(t) == 2; classical RK4, n_steps steps."""
        assert len(t) == 2
        t0, t1 = float(t[0]), float(t[1])
        h = (t1 - t0) / n_steps
        y = tuple(s.detach() for s in y0)

        def f(tt, s):
            out = func(tt, tuple(v.detach().clone() for v in s))
            return tuple(o.detach() for o in out)

        def axpy(s, a, k):
            return tuple(v + a * kv for v, kv in zip(s, k))

        for k in range(n_steps):
            tk = t0 + k * h
            k1 = f(tk, y)
            k2 = f(tk + 0.5 * h, axpy(y, 0.5 * h, k1))
            k3 = f(tk + 0.5 * h, axpy(y, 0.5 * h, k2))
            k4 = f(tk + h, axpy(y, h, k3))
            y = tuple(v + (h / 6.0) * (a + 2 * b + 2 * c + d) for v, a, b, c, d in zip(y, k1, k2, k3, k4))
        return tuple(torch.stack([s0.detach(), s1]) for s0, s1 in zip(y0, y))

    return rk4_fixed


def main():
    sys.path.insert(0, HERE)
    import torch
    import ref_import
    ref_import.import_reference(fp64=False)
    import flow_models.ffjord.layers.cnf as cnf_mod
    import ffjord_oracle as FO

    rng = np.random.RandomState(11)
    x64 = np.concatenate([0.9 * rng.randn(160, 10), 2.0 * rng.randn(32, 10)])
    for name, dt in (("f32", torch.float32), ("f64", torch.float64)):
        nf = torch.load(os.path.join(REF, "pretrained_models", "dsprites", "ffjord", "torch_nf_model.pt"),
                        weights_only=False, map_location="cpu")
        nf.eval()
        if dt == torch.float64:
            nf = nf.double()
        flat, d, hidden, n_blocks, t_end = FO.flatten_state_dict(nf.state_dict())
        x = torch.from_numpy(x64).to(dt)
        out = {}
        # (1) vector field + divergence straight from the reference ODEfunc (no integrator)
        ts = [0.0, 0.3125, 1.0]
        vf_f, vf_div = [], []
        for b, cnf in enumerate(nf.model.chain):
            cnf.odefunc.before_odeint()
            for t in ts:
                dy, mdiv = cnf.odefunc(t, (x.clone(), torch.zeros(x.shape[0], 1, dtype=dt)))
                vf_f.append(dy.detach().numpy())
                vf_div.append(-mdiv.detach().numpy()[:, 0])
        out["vf_t"] = np.array(ts)
        out["vf_f"] = np.stack(vf_f).reshape(n_blocks, len(ts), x.shape[0], d)
        out["vf_div"] = np.stack(vf_div).reshape(n_blocks, len(ts), x.shape[0])
        # (2) FFJORD.forward / .backward through the reference classes with the fixed-grid RK4 in place of torchdiffeq
        cnf_mod.odeint = rk4_fixed_factory(N_STEPS)
        z, logdet = nf(x)
        xb = nf.backward(z.detach())
        cnf_mod.odeint = rk4_fixed_factory(2 * N_STEPS)
        z2, logdet2 = nf(x)
        out.update(params=flat.numpy(), x=x.numpy(), z=z.detach().numpy(), logdet=logdet.detach().numpy(),
                   xb=xb.detach().numpy(), d=d, hidden=np.array(hidden), n_blocks=n_blocks, t_end=np.array(t_end),
                   n_steps=N_STEPS,
                   err_halving=np.array([float((z - z2).abs().max()), float((logdet - logdet2).abs().max())]))
        path = os.path.join(OUT, "ffjord_dsprites_%s.npz" % name)
        np.savez_compressed(path, **out)
        print("wrote", path, os.path.getsize(path), "bytes; params", flat.numel(), "d", d, "hidden", hidden, "blocks",
              n_blocks, "T", t_end, "| |z| max %.3f  logdet range [%.3f, %.3f]  round trip %.2e  halving err z %.2e ld %.2e"
              % (float(z.abs().max()), float(logdet.min()), float(logdet.max()), float((xb - x).abs().max()),
                 out["err_halving"][0], out["err_halving"][1]))
        # the restatement against the same vectors, right here
        spec = FO.FfjordSpec(d=d, hidden=hidden, n_blocks=n_blocks, n_steps=N_STEPS, t_end=t_end)
        zo, ldo = FO.integrate(flat, x, spec, want_logdet=True)
        xo, _ = FO.integrate(flat, z.detach(), spec, reverse=True)
        print("   oracle vs reference: z %.2e  logdet %.2e  backward %.2e" % (
            float((zo - z).abs().max()), float((ldo - logdet).abs().max()), float((xo - xb).abs().max())))


if __name__ == "__main__":
    main()
