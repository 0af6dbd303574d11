"""Parameter container mirroring `src/gp_flows/vector_field.py` (VectorField).

Same constructor, same `denses` ModuleList of nn.Linear (so `state_dict()` keys stay
`v.v.denses.{i}.weight|bias`, SURVEY.md section 5 "Checkpoint / resume"), same initialisation
(N(0, 1e-3) weights, zero bias; vector_field.py:63-70).  The arithmetic (`eval`, `jacobian`,
vector_field.py:75-128) is NOT re-implemented in torch here: it lives inside the fused CUDA kernels, which read
the flat parameter vector `[W_1, b_1, ..., W_{L+1}, b_{L+1}]`; the standalone `eval` / `jacobian` / `__call__` of this class
go through `gpflow_potentials` (csrc/gpflow_potentials.cu).
"""
import torch
import torch.nn as nn


class VectorField(nn.Module):
    def __init__(self, dim_in, dim_out, nb_neurons=[], func_name="tanh", bias=True):
        super().__init__()
        if str(func_name) != "tanh":
            raise NotImplementedError(
                "gpflow_b200 kernels implement the tanh potentials only (default res_net.py:146); "
                "activation %r stays on the reference's Python classes" % (func_name,))
        if not bias:
            raise NotImplementedError("gpflow_b200 kernels assume biased layers (VelocityField always uses bias=True)")
        self.dim_in, self.dim_out, self.func_name = dim_in, dim_out, func_name
        self.neurons = [dim_in] + list(nb_neurons) + [dim_out]
        self.denses = nn.ModuleList(
            [nn.Linear(self.neurons[i], self.neurons[i + 1], bias=True) for i in range(len(self.neurons) - 1)])
        for dense in self.denses:
            nn.init.normal_(dense.weight, mean=0.0, std=1e-3)
            nn.init.zeros_(dense.bias)

    def flat_parameters(self) -> torch.Tensor:
        """[W_1, b_1, ..., W_{L+1}, b_{L+1}] as one contiguous vector (differentiable view-free copy)."""
        return torch.cat([p.reshape(-1) for p in self.parameters()])

    def eval(self, x):  # noqa: A003 - reference method name
        from .ops_aux import potentials
        return potentials(self, x, want_jac=False)[0]

    def jacobian(self, x):
        from .ops_aux import potentials
        return potentials(self, x, want_f=False)[1]

    def __call__(self, x):
        return self.eval(x)
