"""TEST INFRASTRUCTURE ONLY (oracle side) -- import the *unmodified* reference.

Makes `/root/reference` (morel-g/GPFlow, pure Python/PyTorch) importable in the build
container so that golden vectors can be generated from the reference's own code
(`oracle/make_golden.py`) and the oracle restatement (`oracle/gpflow_oracle.py`) can be
pinned against it.  The reference imports two packages that are not installed in this
image (`pytorch_lightning`, `torchdiffeq`); neither is used on the
GP-flow hot path, so in-process stub modules are registered before `src.*` is imported.

`/root/reference` does not exist on the GPU box: nothing in the product package, in
`-m gpu` tests, in `smoke()` or in `bench.py` imports this file.
"""
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("GPFLOW_REFERENCE_ROOT", "/root/reference")


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "src", "gp_flows"))


def _install_stubs():
    import torch

    if "pytorch_lightning" not in sys.modules:
        pl = types.ModuleType("pytorch_lightning")

        class LightningModule(torch.nn.Module):
            def save_hyperparameters(self, *a, **k):
                pass

            def log(self, *a, **k):
                pass

        class LightningDataModule:
            def __init__(self, *a, **k):
                pass

        class Trainer:
            def __init__(self, *a, **k):
                pass

        class Callback:
            pass

        class ModelCheckpoint(Callback):
            def __init__(self, *a, **k):
                pass

        class EarlyStopping(Callback):
            def __init__(self, *a, **k):
                pass

        class TensorBoardLogger:
            def __init__(self, *a, **k):
                pass

        pl.LightningModule = LightningModule
        pl.LightningDataModule = LightningDataModule
        pl.Trainer = Trainer
        pl.Callback = Callback
        pl.seed_everything = lambda seed, **k: torch.manual_seed(seed)
        loggers = types.ModuleType("pytorch_lightning.loggers")
        loggers.TensorBoardLogger = TensorBoardLogger
        tb = types.ModuleType("pytorch_lightning.loggers.tensorboard")
        tb.TensorBoardLogger = TensorBoardLogger
        cbs = types.ModuleType("pytorch_lightning.callbacks")
        cbs.Callback = Callback
        cbs.ModelCheckpoint = ModelCheckpoint
        es = types.ModuleType("pytorch_lightning.callbacks.early_stopping")
        es.EarlyStopping = EarlyStopping
        pl.loggers = loggers
        pl.callbacks = cbs
        sys.modules["pytorch_lightning"] = pl
        sys.modules["pytorch_lightning.loggers"] = loggers
        sys.modules["pytorch_lightning.loggers.tensorboard"] = tb
        sys.modules["pytorch_lightning.callbacks"] = cbs
        sys.modules["pytorch_lightning.callbacks.early_stopping"] = es
    if "torchdiffeq" not in sys.modules:
        try:
            import torchdiffeq  # noqa: F401
        except Exception:
            td = types.ModuleType("torchdiffeq")
            td.odeint_adjoint = None
            td.odeint = None
            sys.modules["torchdiffeq"] = td


def _select_fp64():
    """The reference's fp64 mode is a source-level switch: edit src/precision.py:7-8 to
    `torch_float_precision = torch.double` and call torch.set_default_dtype (main.py:83).
    /root/reference is read-only, so the edited module is provided in-process instead: same
    three names, same values the edited file would define."""
    import numpy as np
    import torch

    prec = types.ModuleType("src.precision")
    prec.torch_float_precision = torch.double
    prec.np_float_precision = np.float64
    prec.eps_precision = np.finfo(np.float64).eps
    sys.modules["src.precision"] = prec
    torch.set_default_dtype(torch.double)


def import_reference(fp64: bool = False):
    """Return the reference's `src.gp_flows.res_net` module (unmodified code).
    fp64=True must be chosen before the first import in a process (like the source switch)."""
    if not reference_available():
        raise RuntimeError("reference tree not present at %s" % REFERENCE_ROOT)
    _install_stubs()
    if fp64:
        if "src.gp_flows.res_net" in sys.modules and "src.precision" in sys.modules and \
                sys.modules["src.precision"].eps_precision > 1e-10:
            raise RuntimeError("reference already imported in fp32 mode in this process")
        _select_fp64()
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    import importlib

    return importlib.import_module("src.gp_flows.res_net")
