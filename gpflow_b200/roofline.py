"""Algorithmic flop count of the GP-flow path -- the numerator of `bench.py`'s roofline object.

Pruned count per velocity evaluation (FMA = 2 flops; only the `(i >= n, j >= n)` entries that enter the velocity,
`velocity_field.py:20-79`; tanh / erf counted separately):

    F_val = sum_l 2 n_l n_{l+1} + 2 H_L R                R  = sum_{n<nb} (d - n)
    F_jac = H_1 d + sum_{l>=2} (2 H_{l-1} H_l d + H_l d) + 2 H_L R2      R2 = sum_{n<nb} (d - n)^2
    F_asm = 6 R2

forward = 4 N F_eval, train step = 3 x forward.  The Euler penalization (`res_net.py:319-358`) adds the NOMINAL
increment of 10 N F_eval forward flops (2 evaluation points per step, each = base evaluation + 2 directional
derivatives at about 2 F_eval apiece), x 3 for the train step.  The kernels may do fewer flops than this count (the
generic kernel evaluates the last layer in contracted form); the count is the survey's definition, not the kernels'.
"""


def flops_per_eval(d, nb, hidden):
    H = list(hidden)
    n = [d + 1] + H
    R = sum(d - k for k in range(nb))
    R2 = sum((d - k) ** 2 for k in range(nb))
    f_val = sum(2 * n[l] * n[l + 1] for l in range(len(n) - 1)) + 2 * H[-1] * R
    f_jac = H[0] * d + sum(2 * H[l - 1] * H[l] * d + H[l] * d for l in range(1, len(H))) + 2 * H[-1] * R2
    f_asm = 6 * R2
    return float(f_val + f_jac + f_asm)


def algorithmic_flops_per_sample(d, nb, hidden, n_steps, train=True, euler=False):
    fwd = (4 + (10 if euler else 0)) * n_steps * flops_per_eval(d, nb, hidden)
    return 3.0 * fwd if train else fwd


def executed_flops_per_sample_tc(d, nb, hidden, n_steps, train=True):
    """Flops the chunked tensor-core path (gpflow_tc*.cuh) EXECUTES per sample, fp32-equivalent (each 3xTF32 product
    counted once), on the zero-padded widths.  Per velocity evaluation, forward: first layer 2 d H_1, hidden layers
    2 (d + 1) H_{l-1} H_l (value row + d tangent rows), last layer in contracted form: F for all d nb rows, D and G for
    the R kept rows, 2 H_L each.  The reverse sweep of an evaluation runs a data GEMM and a weight-gradient GEMM of the
    same size for every forward GEMM (2x), and the train step recomputes the forward evaluation before reversing it:
    train = 4 N (forward + recompute + 2 forward) = 4 x forward."""
    H = [(h + 63) // 64 * 64 for h in hidden]
    R = sum(d - k for k in range(nb))
    f = 2.0 * d * H[0] + sum(2.0 * (d + 1) * H[l - 1] * H[l] for l in range(1, len(H))) + 2.0 * H[-1] * (d * nb + 2 * R)
    fwd = 4 * n_steps * f
    return 4.0 * fwd if train else fwd


def executed_flops_per_sample_sf2r(h1, h2, n_steps):
    """Flops the 2-D stream-function kernel `gpflow_sf2r_kernel` (gpflow_sf2r*.cuh) EXECUTES per sample and train step
    (FMA = 2, tanh / erf not counted, like the algorithmic count).  It obtains grad psi in reverse mode, so a forward
    evaluation has two H1 x H2 contractions (a2, z = W2^T ws) instead of the three the forward-mode count assumes, and
    its reverse sweep three (tb, hb, z again) plus a rank-2 update of dW2 instead of three plus a rank-3 update:
        F_fwd = 6 H1 + 2 H1 H2 + H2 + 2 H1 H2 + 2 H2 + 3 H2 + 2 H1 + 4 H1 + 12
        F_rev = 5 H1 + 3 (2 H1 H2) + 10 H2 + 14 H1 + 4 H2 (H1 + 1) + 11 H1 + 40
    (layer 1, a2 + bias, z, psi, sp2 / ws, sp1 / q, grad psi, closed form | c, the three contractions, per-unit adjoints of
    layer 2 and layer 1, dW2 with its bias column, dW1 / db1, closed-form adjoint).  H = [15, 15]: 4 132 per evaluation
    against the algorithmic 5 112 (0.81)."""
    f_fwd = 6 * h1 + 2 * h1 * h2 + h2 + 2 * h1 * h2 + 2 * h2 + 3 * h2 + 2 * h1 + 4 * h1 + 12
    f_rev = 5 * h1 + 3 * (2 * h1 * h2) + 10 * h2 + 14 * h1 + 4 * h2 * (h1 + 1) + 11 * h1 + 40
    return 4.0 * n_steps * (f_fwd + f_rev)
