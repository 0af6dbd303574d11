"""The INTEGRATION.md section 1 hook, applied to the UNMODIFIED reference imported in this container (not gpu; skipped on
the GPU box where /root/reference does not exist): `gpflow_b200.integration.install()` rebinds `src.models.get_gp_flow`,
the single place the reference builds its GP flow (src/models.py:36-45), and the reference's own `Data` dictionaries
(src/data.py:69-81, default_params/toy_params.py:93-104, default_params/latent_params.py:195-205) must then yield the CUDA
mirror -- not the fall-through -- with the reference's constants (`Case.continuous_time == "continuous time"`)."""
import importlib
import logging

import pytest
import torch

import ref_import

pytestmark = pytest.mark.skipif(not ref_import.reference_available(), reason="reference tree not present")


@pytest.fixture()
def models():
    ref_import.import_reference()
    m = importlib.import_module("src.models")
    import gpflow_b200.integration as gi
    gi.uninstall(m)
    yield m
    gi.uninstall(m)


def _call(m, data):
    return m.get_gp_flow(data.dim, data.velocity_dict, data.nb_layers_gp, data.nb_blocks, data.T_final, data.ode_params,
                         data.euler_dict, data.euler_spectral_velocity_dict)


def _toy(use_euler=False):
    from src.case import Case
    from default_params.toy_params import get_toy_params
    return get_toy_params(Case.eight_gaussians, Case.bnaf, Case.train_gp, model_path="pretrained_models/eight_gaussians/bnaf",
                          use_euler=use_euler, euler_case=Case.penalization)


def test_hook_returns_cuda_mirror_for_reference_default_params(models, caplog):
    import gpflow_b200 as G
    import gpflow_b200.integration as gi
    ref_cls = importlib.import_module("src.gp_flows.res_net").ResNet
    data = _toy()
    assert isinstance(_call(models, data), ref_cls)          # before the hook: the reference's class
    hook = gi.install(models)
    assert models.get_gp_flow is hook and gi.install(models) is hook      # idempotent
    with caplog.at_level(logging.INFO, logger="gpflow_b200.integration"):
        flow = _call(models, data)
    assert isinstance(flow, G.ResNet) and not isinstance(flow, ref_cls)
    assert "CUDA path" in caplog.text
    assert flow.nb_layers == data.nb_layers_gp == 15 and flow.v.nb_div_blocks == 1
    assert [tuple(l.weight.shape) for l in flow.v.v.denses] == [(15, 3), (15, 15), (2, 15)]
    # state_dict keys interchange with the reference's module (checkpoints, training.py:48-55)
    ref_flow = importlib.import_module("src.gp_flows.res_net").get_gp_flow(
        data.dim, data.velocity_dict, data.nb_layers_gp, data.nb_blocks, data.T_final, data.ode_params, data.euler_dict,
        data.euler_spectral_velocity_dict)
    assert list(flow.state_dict().keys()) == list(ref_flow.state_dict().keys())
    flow.load_state_dict(ref_flow.state_dict())
    # Euler penalization: the coefficient the callback mutates (gp_flow_callback.py:13-26)
    flow_e = _call(models, _toy(use_euler=True))
    assert isinstance(flow_e, G.ResNet) and flow_e.euler_coef_penalization == 5e-4


def test_hook_accepts_the_reference_constants(models):
    """An explicit `time_state = Case.continuous_time` ("continuous time", src/case.py:4), `Case.RK4`, `Case.standard`,
    `Case.tanh` must stay on the CUDA path; nb_blocks = 2 gives the Compose mirror."""
    import gpflow_b200 as G
    import gpflow_b200.integration as gi
    from gpflow_b200.res_net import Compose
    from src.case import Case
    gi.install(models)
    data = _toy()
    data.velocity_dict = dict(data.velocity_dict, time_state=Case.continuous_time, activ_func=Case.tanh, nb_div_blocks=1)
    data.ode_params = dict(data.ode_params, method=Case.RK4, solve_method=Case.standard)
    flow = _call(models, data)
    assert isinstance(flow, G.ResNet) and flow.v.time_state == Case.continuous_time
    data.nb_blocks = 2
    comp = _call(models, data)
    assert isinstance(comp, Compose) and len(comp.layers) == 2 and all(isinstance(l, G.ResNet) for l in comp.layers)


def test_hook_falls_through_loudly(models, caplog):
    """Options outside the CUDA path keep the reference's classes and say so in the log."""
    import gpflow_b200.integration as gi
    from src.case import Case
    ref_cls = importlib.import_module("src.gp_flows.res_net").ResNet
    gi.install(models)
    for change in (dict(time_state=Case.stationary), dict(activ_func=Case.relu), dict(incompressible=False)):
        data = _toy()
        data.velocity_dict = dict(data.velocity_dict, **change)
        with caplog.at_level(logging.WARNING, logger="gpflow_b200.integration"):
            caplog.clear()
            flow = _call(models, data)
        assert isinstance(flow, ref_cls), change
        assert "keeping the reference" in caplog.text, change
    data = _toy()
    data.ode_params = dict(data.ode_params, method=Case.euler)
    with caplog.at_level(logging.WARNING, logger="gpflow_b200.integration"):
        caplog.clear()
        assert isinstance(_call(models, data), ref_cls)
        assert "keeping the reference" in caplog.text


def test_latent_params_stay_on_cuda_path(models):
    """C3 hyper-parameters (default_params/latent_params.py:195-205: d = 10, nb_div_blocks = 9, [50, 50, 50])."""
    import gpflow_b200 as G
    import gpflow_b200.integration as gi
    gi.install(models)
    vd = {"nb_neurons": [50, 50, 50], "nb_div_blocks": 9}
    flow = models.get_gp_flow(10, vd, 15, 1, 1.0, {"method": "RK4"}, {"use_euler": False}, None)
    assert isinstance(flow, G.ResNet) and flow.v.nb_div_blocks == 9
    assert sum(p.numel() for p in flow.parameters()) == 10290
    assert flow.v.bound == 1 - float(torch.finfo(torch.float32).eps)
