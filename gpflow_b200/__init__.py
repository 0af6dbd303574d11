"""gpflow_b200 -- B200-native (sm_100a) GP-flow (Gaussian-preserving flow) hot path of morel-g/GPFlow.

Host side: a mirror of the reference's `src/gp_flows` module interface (`get_gp_flow`, `ResNet`,
`VelocityField`, `VectorField`, `Network`-level `gp_flow` / `loss_gp_flow` / fused train step) on top of
a C-ABI CUDA library (`include/gpflow_b200.h`, `gpflow_b200/csrc`).  There is no CPU fallback: every op
raises if the CUDA library is missing or the tensors are not on a CUDA device.
"""
from . import _lib  # noqa: F401
from .ops import FlowConfig, flow_forward, flow_backward, fused_train_step, kernel_variant  # noqa: F401
from .res_net import get_gp_flow, ResNet  # noqa: F401
from .velocity_field import VelocityField  # noqa: F401
from .vector_field import VectorField  # noqa: F401
from .network import GPFlowNetwork  # noqa: F401
from .host_stream import HostBatchStream  # noqa: F401
from .optim import FlatAdam  # noqa: F401
from .bnaf import BNAFMap  # noqa: F401
from .ffjord import FFJORDMap, gaussian_noise_dataset  # noqa: F401

__all__ = ["FlowConfig", "flow_forward", "flow_backward", "fused_train_step", "kernel_variant", "get_gp_flow",
           "ResNet", "VelocityField", "VectorField", "GPFlowNetwork", "HostBatchStream", "FlatAdam", "BNAFMap", "FFJORDMap",
           "gaussian_noise_dataset"]
