"""The hook that puts this package behind the reference's `main.py --opt_type train_gp`.

The reference builds its GP flow in exactly one place: `src/models.py:36-45` calls `get_gp_flow(dim, velocity_dict,
nb_layers_gp, nb_blocks, T_final, ode_params, euler_dict, euler_spectral_velocity_dict)`, a name `src/models.py:13`
imports from `src/gp_flows/res_net.py:14-67`.  `install()` rebinds that name to a wrapper which returns the CUDA mirror
(`gpflow_b200.get_gp_flow`, same signature, same returned interface) and falls through -- LOUDLY, through `logging` --
to the reference's own classes for the options the kernels do not implement (spectral Euler, relu / log_cosh,
stationary or compressible velocities, non-RK4 solvers).  The fall-through is the untouched reference
code, not a CPU version of the kernels: a module returned by the mirror raises on CPU tensors.

    import gpflow_b200.integration as gi; gi.install()      # before `python main.py ...` builds the Network
"""
import logging
import sys

log = logging.getLogger("gpflow_b200.integration")


def make_hook(reference_get_gp_flow, dtype=None):
    """Wrap the reference's `get_gp_flow`; `dtype` defaults to the reference's precision mode (src/precision.py:7-8)."""
    from .res_net import get_gp_flow as cuda_get_gp_flow

    def get_gp_flow(dim, velocity_dict, nb_layers_gp, nb_blocks=1, T_final=1.0, ode_params={"method": "RK4"},
                    euler_dict={"use_euler": False}, euler_spectral_velocity_dict=None):
        import torch
        dt = dtype
        if dt is None:
            prec = sys.modules.get("src.precision")
            dt = getattr(prec, "torch_float_precision", torch.float32)
        try:
            flow = cuda_get_gp_flow(dim, velocity_dict, nb_layers_gp, nb_blocks, T_final, ode_params, euler_dict,
                                    euler_spectral_velocity_dict, dtype=dt)
            log.info("gpflow_b200: GP flow on the CUDA path (dim=%d, nb_layers=%d, nb_blocks=%d)", dim, nb_layers_gp, nb_blocks)
            return flow
        except NotImplementedError as e:
            log.warning("gpflow_b200: option outside the CUDA path, keeping the reference's Python classes: %s", e)
            return reference_get_gp_flow(dim, velocity_dict, nb_layers_gp, nb_blocks, T_final, ode_params, euler_dict,
                                         euler_spectral_velocity_dict)

    get_gp_flow.__wrapped__ = reference_get_gp_flow
    get_gp_flow.gpflow_b200_hook = True
    return get_gp_flow


def install(models_module=None, dtype=None):
    """Rebind `src.models.get_gp_flow` (and `src.gp_flows.res_net.get_gp_flow` for direct importers) to the hook.
    Returns the hook.  Idempotent."""
    if models_module is None:
        import importlib
        models_module = importlib.import_module("src.models")
    current = models_module.get_gp_flow
    if getattr(current, "gpflow_b200_hook", False):
        return current
    hook = make_hook(current, dtype=dtype)
    models_module.get_gp_flow = hook
    return hook


def uninstall(models_module=None):
    if models_module is None:
        models_module = sys.modules.get("src.models")
    if models_module is None:
        return
    current = models_module.get_gp_flow
    if getattr(current, "gpflow_b200_hook", False):
        models_module.get_gp_flow = current.__wrapped__
