"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol that
include/gpflow_b200.h declares, validates descriptors, and the host mirror keeps the reference's interface.
No compute call is made here (no GPU in the build container)."""
import ctypes
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from gpflow_b200 import build, _lib
    build.build()
    return _lib.load()


def test_header_symbols_are_exported(lib):
    hdr = open(os.path.join(ROOT, "include", "gpflow_b200.h")).read()
    declared = set(re.findall(r"\b(gpflow_[a-z0-9_]+)\s*\(", hdr))
    from gpflow_b200 import _lib
    assert declared == set(_lib.EXPORTS), declared ^ set(_lib.EXPORTS)
    for sym in declared:
        assert getattr(lib, sym) is not None


def test_desc_struct_layout_matches_header():
    from gpflow_b200._lib import GpFlowDesc
    # 4 + 4 + 4 + 4 + 4 + 8*4 + 4 + 4 + 4 + 4 = 68 -> padded to 72, then 5 doubles
    assert ctypes.sizeof(GpFlowDesc) == 72 + 5 * 8
    assert GpFlowDesc.dt.offset == 72 and GpFlowDesc.loss_scale.offset == 104


def test_param_count_and_validation(lib):
    from gpflow_b200 import ops, FlowConfig, _lib
    for cfg, want in [(FlowConfig(2, 1, (15, 15), 15), 332), (FlowConfig(10, 9, (50, 50, 50), 15), 10290),
                      (FlowConfig(64, 63, (256, 256), 15), 1118912)]:
        d = ops.make_desc(cfg, torch.float32)
        assert lib.gpflow_param_count(ctypes.byref(d)) == want == cfg.n_params
    bad = ops.make_desc(FlowConfig(2, 1, (15, 15), 15), torch.float32)
    bad.nb = 3
    assert lib.gpflow_param_count(ctypes.byref(bad)) == -1
    n = ctypes.c_size_t(0)
    assert lib.gpflow_query_workspace(ctypes.byref(bad), 10, 0, ctypes.byref(n)) == 1
    bad.nb = 1
    bad.abi_version = 99
    assert lib.gpflow_query_workspace(ctypes.byref(bad), 10, 0, ctypes.byref(n)) == 1
    assert b"descriptor" in lib.gpflow_error_string(1)
    assert lib.gpflow_abi_version() == _lib.GPFLOW_ABI_VERSION


def test_workspace_query_and_variant_names(lib):
    from gpflow_b200 import ops, FlowConfig, kernel_variant, _lib
    c4 = FlowConfig(2, 1, (15, 15), 15)
    assert "gpflow_sf2r_kernel<f32,d=2,nb=1,H=[15,15],train>" == kernel_variant(c4)
    assert "gpflow_small_kernel<f32,d=2,nb=1,H=[15,15],train>" == kernel_variant(c4, force_v1=True)
    assert "gpflow_sf2e_kernel<f32,d=2,nb=1,H=[15,15],train,euler>" == kernel_variant(c4, euler_coef=0.3)
    assert "gpflow_small_kernel" in kernel_variant(FlowConfig(2, 1, (15, 15), 600), euler_coef=0.3)   # buffers would not fit
    assert "gpflow_small_kernel" in kernel_variant(c4, torch.float64, euler_coef=0.3)
    assert "gpflow_sf2r_fwd_kernel<f32,d=2,nb=1,H=[15,15],fwd>" == kernel_variant(c4, op=_lib.OP_FORWARD, lagrangian=True)
    assert "gpflow_small_kernel<f32,d=2,nb=1,H=[15,15],fwd>" == kernel_variant(c4, op=_lib.OP_FORWARD, force_v1=True)
    assert "gpflow_small_kernel<f64" in kernel_variant(c4, torch.float64, op=_lib.OP_FORWARD)
    # narrower 2-D nets: the [15,15] kernels on zero-padded parameters (gpflow_api.cu: pad2d_match)
    narrow = FlowConfig(2, 1, (10, 12), 15)
    assert "gpflow_sf2r_kernel<f32,d=2,nb=1,H=[15,15],train> (H=[10,12] zero-padded)" == kernel_variant(narrow)
    assert "gpflow_sf2e_kernel" in kernel_variant(narrow, euler_coef=0.3) and "zero-padded" in kernel_variant(narrow, euler_coef=0.3)
    assert "gpflow_generic_kernel" in kernel_variant(narrow, torch.float64)
    assert "zero-padded" not in kernel_variant(FlowConfig(2, 1, (16, 10), 15))      # wider than 15: tc / generic
    assert "generic" in kernel_variant(c4, force_generic=True)
    assert "generic" in kernel_variant(FlowConfig(10, 9, (50, 50, 50), 15))
    assert "euler" in kernel_variant(c4, euler_coef=0.01)
    assert "euler" not in kernel_variant(c4, euler_coef=1e-9)     # coef <= 2 eps disables the penalty (res_net.py:299)
    n = ctypes.c_size_t(0)
    d = ops.make_desc(c4, torch.float32, wrap=True)
    assert lib.gpflow_query_workspace(ctypes.byref(d), 1 << 24, _lib.OP_TRAIN_STEP, ctypes.byref(n)) == 0
    assert 0 < n.value < 1200 * (1 << 20)    # persistent grid: the activation stash (<= 160 SMs x 12 warps x 60 x 9 KB) does not scale with B
    n_small = ctypes.c_size_t(0)
    assert lib.gpflow_query_workspace(ctypes.byref(d), 1024, _lib.OP_TRAIN_STEP, ctypes.byref(n_small)) == 0
    assert 0 < n_small.value < (1 << 25)
    big = ops.make_desc(FlowConfig(64, 63, (256, 256), 15), torch.float32, wrap=True)
    assert lib.gpflow_query_workspace(ctypes.byref(big), 1 << 14, _lib.OP_TRAIN_STEP, ctypes.byref(n)) == 0
    assert n.value < (17 << 30)    # <= 16 GiB of per-warp scratch + gradient partials (1.1 M parameters per warp) + finalize buffers


def test_module_mirror_keeps_reference_interface():
    import gpflow_b200 as G
    flow = G.get_gp_flow(2, {"nb_neurons": [15, 15]}, 15,
                         euler_dict={"use_euler": True, "case": "penalization", "coef_penalization": 5e-4})
    assert list(flow.state_dict().keys()) == ["v.v.denses.%d.%s" % (i, w) for i in range(3) for w in ("weight", "bias")]
    assert flow.nb_layers == 15 and abs(flow.dt - 1 / 15) < 1e-15 and flow.euler_coef_penalization == 5e-4
    assert flow.v.nb_div_blocks == 1 and abs(flow.v.bound - (1 - 1.1920929e-07)) < 1e-12
    flow.set_nb_layers_eval(30)
    assert flow.nb_layers_eval == 30 and abs(flow.dt_eval - 1 / 30) < 1e-15
    assert float(flow.regularization_coef()) == 0.0
    # reference initialisation: N(0, 1e-3) weights, zero bias (vector_field.py:63-70)
    assert all(float(dn.bias.abs().max()) == 0.0 for dn in flow.v.v.denses)
    assert 1e-4 < float(flow.v.v.denses[1].weight.std()) < 3e-3
    comp = G.get_gp_flow(2, {"nb_neurons": [15, 15]}, 15, nb_blocks=2)
    assert len(comp.layers) == 2
    with pytest.raises(NotImplementedError):
        G.get_gp_flow(2, {"nb_neurons": [15], "activ_func": "relu"}, 15)
    with pytest.raises(NotImplementedError):
        G.get_gp_flow(2, {"nb_neurons": [15]}, 15, euler_dict={"use_euler": True, "case": "spectral_method"})


def test_no_cpu_fallback():
    import gpflow_b200 as G
    cfg = G.FlowConfig(2, 1, (15, 15), 15)
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        G.flow_forward(cfg, torch.zeros(cfg.n_params), torch.zeros(4, 2))
    flow = G.get_gp_flow(2, {"nb_neurons": [15, 15]}, 15)
    with pytest.raises(RuntimeError, match="CUDA tensor"):
        flow(torch.zeros(4, 2))


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "gpflow_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dp, f)).read()
                assert "gpflow_oracle" not in src and "ref_import" not in src, f


def test_shipped_library_is_a_default_build_of_these_sources(lib):
    """The in-tree .so is what the GPU box runs: it must come from the committed sources with the default flags (an
    experiment build made with GPFLOW_NVCC_EXTRA is marked in the hash file and rejected here and by smoke())."""
    from gpflow_b200 import build as b
    assert b.built_hash() is not None and "+extra" not in b.built_hash()
    assert b.built_hash() == b.source_hash()


def test_draw_probes_keeps_the_reference_rng_stream():
    """res_net.py:330-333 draws z then w with torch.randn(1, d) once per step on the default generator; the mirror must
    consume the generator identically (the C2 training replay depends on it) whatever staging it uses for the upload."""
    import gpflow_b200 as G
    flow = G.get_gp_flow(3, {"nb_neurons": [5, 5], "nb_div_blocks": 2}, 7,
                         euler_dict={"use_euler": True, "case": "penalization", "coef_penalization": 5e-4})
    torch.manual_seed(11)
    got = flow.draw_probes(7, torch.zeros(1, dtype=torch.float64))
    after = torch.randn(1)
    torch.manual_seed(11)
    rows = []
    for _ in range(7):
        z = torch.randn(1, 3)
        w = torch.randn(1, 3)
        rows.append(torch.stack([z[0], w[0]]))
    want = torch.stack(rows).double()
    assert got.dtype == torch.float64 and torch.equal(got, want)
    assert torch.equal(after, torch.randn(1))          # and leaves the generator in the same state
