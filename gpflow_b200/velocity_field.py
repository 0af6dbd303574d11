"""Mirror of `src/gp_flows/velocity_field.py` (VelocityField): the divergence-free, wall-tangent velocity.

Holds the potentials' parameters (`self.v`, a VectorField) and the boundary `bound = 1 - eps`
(velocity_field.py:96).  `__call__(x, t)` evaluates the velocity on the GPU through the C ABI; the training /
transport path never calls it (the fused kernels evaluate the velocity internally).
"""
import numpy as np
import torch

from .vector_field import VectorField


# src/case.py:3-4 -- the reference's constants are the strings below ("continuous time" has a SPACE); the underscore
# spelling is accepted as an alias because round-1 callers of this mirror used it.
CONTINUOUS_TIME = "continuous time"
STATIONARY = "stationary"
_CONTINUOUS_ALIASES = (CONTINUOUS_TIME, "continuous_time")


class VelocityField(torch.nn.Module):
    def __init__(self, dim, nb_neurons, activ_func="tanh", time_state=CONTINUOUS_TIME, incompressible=True,
                 nb_div_blocks=1, dtype=torch.float32):
        super().__init__()
        if str(time_state) not in _CONTINUOUS_ALIASES:
            raise NotImplementedError("gpflow_b200: only Case.continuous_time (%r) is on the CUDA path, got %r"
                                      % (CONTINUOUS_TIME, time_state))
        time_state = CONTINUOUS_TIME
        if not incompressible:
            raise NotImplementedError("gpflow_b200: only the incompressible (Gaussian-preserving) velocity is on the CUDA path")
        self.time_state, self.incompressible = time_state, incompressible
        self.dim = self.dim_in = self.dim_out = dim
        self.nb_div_blocks = nb_div_blocks
        self.bound = 1 - float(np.finfo(np.float64 if dtype == torch.float64 else np.float32).eps)
        self.v = VectorField(dim + 1, dim * nb_div_blocks, list(nb_neurons), func_name=activ_func)
        if dtype == torch.float64:
            self.v.double()

    def flat_parameters(self):
        return self.v.flat_parameters()

    def __call__(self, x, t=0):
        from .ops_aux import velocity
        return velocity(self, x, t)
