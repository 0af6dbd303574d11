"""ctypes binding of the C ABI declared in include/gpflow_b200.h (the drop-in boundary).

The shared library is built in-tree (gpflow_b200/lib/libgpflow_b200.so) by `gpflow_b200.build.build()` /
`__graft_entry__.build()`; it links only against the CUDA runtime -- no torch symbols.
"""
import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# GPFLOW_B200_LIB: developer override to A/B-test an alternative build of the same library (still in-tree, still CUDA)
LIB_PATH = os.environ.get("GPFLOW_B200_LIB") or os.path.join(_HERE, "lib", "libgpflow_b200.so")

GPFLOW_ABI_VERSION = 1
ABI_VERSION = GPFLOW_ABI_VERSION
ERR_BAD_DESC = 1       # include/gpflow_b200.h: GPFLOW_ERR_BAD_DESC
ERR_UNSUPPORTED = 2    # GPFLOW_ERR_UNSUPPORTED
ERR_NULL_POINTER = 3   # GPFLOW_ERR_NULL_POINTER
ERR_STATIC_SCRATCH = 7  # GPFLOW_ERR_STATIC_SCRATCH
ERR_COLLECTIVE = 8  # GPFLOW_ERR_COLLECTIVE
STATIC_SCRATCH_BYTES = 32 << 20  # GPFLOW_STATIC_SCRATCH_BYTES
GPFLOW_MAX_HIDDEN = 8
F32, F64 = 0, 1
OP_FORWARD, OP_TRAIN_STEP, OP_BACKWARD = 0, 1, 2
FLAG_WRAP_GAUSS, FLAG_LAGRANGIAN, FLAG_EULER_PEN, FLAG_REVERSE, FLAG_FORCE_GENERIC, FLAG_FORCE_V1 = 0x1, 0x2, 0x4, 0x8, 0x10, 0x20
FLAG_FORCE_TC = 0x40
SCALAR_OT, SCALAR_REG, SCALAR_LAGR, NUM_SCALARS = 0, 1, 2, 3

EXPORTS = [
    "gpflow_param_count", "gpflow_query_workspace", "gpflow_forward", "gpflow_train_step", "gpflow_backward",
    "gpflow_velocity", "gpflow_velocity_static", "gpflow_potentials", "gpflow_kernel_variant", "gpflow_launch_count", "gpflow_tc_launches_total", "gpflow_tc_slab_samples", "gpflow_fp32_fma_burn", "gpflow_tf32_mma_burn", "gpflow_clock_probe",
    "gpflow_adam_step", "gpflow_bnaf_param_count", "gpflow_bnaf_forward", "gpflow_ffjord_param_count",
    "gpflow_ffjord_integrate", "gpflow_error_string",
    "gpflow_abi_version",
    "gpflow_comm_unique_id", "gpflow_comm_create", "gpflow_allreduce_grads", "gpflow_comm_destroy",
]


class GpFlowBnafDesc(ctypes.Structure):
    """include/gpflow_b200.h: GpFlowBnafDesc."""
    _fields_ = [("abi_version", ctypes.c_int32), ("dtype", ctypes.c_int32), ("dim", ctypes.c_int32),
                ("hidden_per_dim", ctypes.c_int32), ("n_layers", ctypes.c_int32), ("reserved0", ctypes.c_int32)]


FFJORD_MAX_HIDDEN, FFJORD_MAX_BLOCKS = 7, 8


class GpFlowFfjordDesc(ctypes.Structure):
    """include/gpflow_b200.h: GpFlowFfjordDesc."""
    _fields_ = [("abi_version", ctypes.c_int32), ("dtype", ctypes.c_int32), ("dim", ctypes.c_int32),
                ("n_hidden", ctypes.c_int32), ("hidden", ctypes.c_int32 * FFJORD_MAX_HIDDEN), ("n_blocks", ctypes.c_int32),
                ("n_steps", ctypes.c_int32), ("reverse", ctypes.c_int32), ("t_end", ctypes.c_double * FFJORD_MAX_BLOCKS)]


class GpFlowDesc(ctypes.Structure):
    _fields_ = [
        ("abi_version", ctypes.c_int32),
        ("dtype", ctypes.c_int32),
        ("d", ctypes.c_int32),
        ("nb", ctypes.c_int32),
        ("n_hidden", ctypes.c_int32),
        ("hidden", ctypes.c_int32 * GPFLOW_MAX_HIDDEN),
        ("n_steps", ctypes.c_int32),
        ("n_steps_nominal", ctypes.c_int32),
        ("flags", ctypes.c_uint32),
        ("reserved0", ctypes.c_int32),
        ("dt", ctypes.c_double),
        ("dt_constraint", ctypes.c_double),
        ("sigma", ctypes.c_double),
        ("euler_coef", ctypes.c_double),
        ("loss_scale", ctypes.c_double),
    ]


_lib = None


class GpFlowLibraryError(RuntimeError):
    pass


def load():
    """Load the CUDA library; fail loudly if it has not been built (there is no fallback path)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise GpFlowLibraryError(
            "gpflow_b200 CUDA library not built: %s is missing. Build it with "
            "`python -c 'import __graft_entry__ as g; g.build()'` (needs nvcc). There is no CPU fallback." % LIB_PATH)
    import torch  # noqa: F401  -- loads the CUDA runtime (libcudart.so.12) this library is dynamically linked against
    lib = ctypes.CDLL(LIB_PATH)
    P, D = ctypes.c_void_p, ctypes.POINTER(GpFlowDesc)
    lib.gpflow_param_count.restype = ctypes.c_int64
    lib.gpflow_param_count.argtypes = [D]
    lib.gpflow_query_workspace.restype = ctypes.c_int
    lib.gpflow_query_workspace.argtypes = [D, ctypes.c_int64, ctypes.c_int, ctypes.POINTER(ctypes.c_size_t)]
    lib.gpflow_forward.restype = ctypes.c_int
    lib.gpflow_forward.argtypes = [D, ctypes.c_int64, P, P, P, P, P, P, P, P, ctypes.c_size_t, P]
    lib.gpflow_train_step.restype = ctypes.c_int
    lib.gpflow_train_step.argtypes = [D, ctypes.c_int64, P, P, P, P, P, P, P, P, P, ctypes.c_size_t, P]
    lib.gpflow_backward.restype = ctypes.c_int
    lib.gpflow_backward.argtypes = [D, ctypes.c_int64, P, P, P, P, P, P, P, P, ctypes.c_size_t, P]
    lib.gpflow_velocity.restype = ctypes.c_int
    lib.gpflow_velocity.argtypes = [D, ctypes.c_int64, P, P, ctypes.c_double, P, P, ctypes.c_size_t, P]
    lib.gpflow_velocity_static.restype = ctypes.c_int
    lib.gpflow_velocity_static.argtypes = [D, ctypes.c_int64, P, P, ctypes.c_double, P, P]
    lib.gpflow_potentials.restype = ctypes.c_int
    lib.gpflow_potentials.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_int32), ctypes.c_int64, P, P, P, P, P]
    lib.gpflow_kernel_variant.restype = ctypes.c_int
    lib.gpflow_kernel_variant.argtypes = [D, ctypes.c_int, ctypes.c_char_p, ctypes.c_size_t]
    lib.gpflow_launch_count.restype = ctypes.c_int
    lib.gpflow_launch_count.argtypes = [D, ctypes.c_int]
    lib.gpflow_tc_launches_total.restype = ctypes.c_int64
    lib.gpflow_tc_launches_total.argtypes = []
    lib.gpflow_tc_slab_samples.restype = ctypes.c_int
    lib.gpflow_tc_slab_samples.argtypes = [D, ctypes.c_int64, ctypes.c_int]
    lib.gpflow_clock_probe.restype = ctypes.c_int
    lib.gpflow_clock_probe.argtypes = [P, P]
    lib.gpflow_comm_unique_id.restype = ctypes.c_int
    lib.gpflow_comm_unique_id.argtypes = [P]
    lib.gpflow_comm_create.restype = ctypes.c_int
    lib.gpflow_comm_create.argtypes = [P, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_void_p)]
    lib.gpflow_allreduce_grads.restype = ctypes.c_int
    lib.gpflow_allreduce_grads.argtypes = [P, ctypes.c_int64, ctypes.c_int, P, P]
    lib.gpflow_comm_destroy.restype = ctypes.c_int
    lib.gpflow_comm_destroy.argtypes = [P]
    BD = ctypes.POINTER(GpFlowBnafDesc)
    lib.gpflow_bnaf_param_count.restype = ctypes.c_int
    lib.gpflow_bnaf_param_count.argtypes = [BD, ctypes.POINTER(ctypes.c_int64)]
    lib.gpflow_bnaf_forward.restype = ctypes.c_int
    lib.gpflow_bnaf_forward.argtypes = [BD, ctypes.c_int64, P, P, P, P, P]
    FD = ctypes.POINTER(GpFlowFfjordDesc)
    lib.gpflow_ffjord_param_count.restype = ctypes.c_int
    lib.gpflow_ffjord_param_count.argtypes = [FD, ctypes.POINTER(ctypes.c_int64)]
    lib.gpflow_ffjord_integrate.restype = ctypes.c_int
    lib.gpflow_ffjord_integrate.argtypes = [FD, ctypes.c_int64, P, P, P, P, P]
    lib.gpflow_adam_step.restype = ctypes.c_int
    lib.gpflow_adam_step.argtypes = [ctypes.c_int, ctypes.c_int64, P, P, P, P] + [ctypes.c_double] * 5 + [ctypes.c_int64, P]
    lib.gpflow_tf32_mma_burn.restype = ctypes.c_int
    lib.gpflow_tf32_mma_burn.argtypes = [ctypes.c_int64, P, ctypes.POINTER(ctypes.c_double), P]
    lib.gpflow_fp32_fma_burn.restype = ctypes.c_int
    lib.gpflow_fp32_fma_burn.argtypes = [ctypes.c_int64, P, ctypes.POINTER(ctypes.c_double), P]
    lib.gpflow_error_string.restype = ctypes.c_char_p
    lib.gpflow_error_string.argtypes = [ctypes.c_int]
    lib.gpflow_abi_version.restype = ctypes.c_int
    lib.gpflow_abi_version.argtypes = []
    if lib.gpflow_abi_version() != GPFLOW_ABI_VERSION:
        raise GpFlowLibraryError("ABI version mismatch between %s and the Python binding" % LIB_PATH)
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        raise RuntimeError("gpflow_b200: " + load().gpflow_error_string(rc).decode())
