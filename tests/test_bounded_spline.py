"""tails=None: the bounded rational-quadratic spline on [0, 1] (utils/splines.py:88-193 called directly from
neural_spline/coupling.py:321-323 and autoregressive.py:106-108) -- the constructor default of the spline layers.
Golden vectors from the unmodified reference: tests/golden/layers_r3.pt (`make_golden.py --r3`).
CPU: the oracle against the goldens; GPU: the CUDA path (kernel tail mode VCNF_TAIL_NONE) against the goldens."""
import os

import pytest
import torch
from torch import nn

import oracle
import vcnf_b200 as nf
from cases import LAYER_CASES_R3
from harness import eval_layer
from parity import assert_close, assert_grads_close, SPLINE

GOLD = os.path.join(os.path.dirname(__file__), "golden", "layers_r3.pt")


def _oracle_layer(name):
    fl, nets = oracle.flows, oracle.nets
    if name == "prq_coupling_bounded_uncond":
        mk = lambda cin, cout: nets.ResidualNet(cin, cout, 16, num_blocks=1)
        return fl._SplineCoupling(fl.alternating_mask(6, True), mk, 8, None, 1.0)
    if name == "prq_cdf_bounded":
        return fl._SharedSpline(5, 8, None, 1.0)
    if name == "marqs_bounded":
        return fl._SplineAutoregressive(6, 16, 8, None, 1.0, 1, False, nn.functional.relu, False)
    return None


@pytest.mark.parametrize("direction", ["forward", "inverse"])
@pytest.mark.parametrize("name", ["prq_coupling_bounded_uncond", "prq_cdf_bounded", "marqs_bounded"])
def test_oracle_bounded_spline_matches_reference(name, direction):
    g = torch.load(GOLD)[name]
    layer = _oracle_layer(name)
    layer.load_state_dict(g["state"], strict=True)
    got = eval_layer(layer, g["x"], direction)
    ref = g[direction]
    assert_close(got["z"], ref["z"], rtol=1e-6, atol=1e-6, what=name + ".z")
    assert_close(got["log_det"], ref["log_det"], rtol=1e-5, atol=1e-5, what=name + ".log_det")
    assert_grads_close({"x": got["gx"], **got["grads"]}, {"x": ref["gx"], **ref["grads"]}, rtol=1e-4, what=name)


@pytest.mark.parametrize("name", sorted(LAYER_CASES_R3))
def test_bounded_spline_state_dict_keys_match_reference(name):
    gold = torch.load(GOLD)[name]
    torch.manual_seed(0)
    layer = LAYER_CASES_R3[name]["build"](nf)
    layer.load_state_dict(gold["state"], strict=True)
    assert list(layer.state_dict().keys()) == list(gold["state"].keys())


@pytest.mark.gpu
@pytest.mark.parametrize("direction", ["inverse", "forward"])
@pytest.mark.parametrize("name", sorted(LAYER_CASES_R3))
def test_bounded_spline_layer_matches_reference_golden(name, direction):
    case, g = LAYER_CASES_R3[name], torch.load(GOLD)[name]
    torch.manual_seed(0)
    layer = case["build"](nf)
    layer.load_state_dict(g["state"])
    layer = layer.cuda()
    before = nf.ops.LAUNCHES
    got = eval_layer(layer, g["x"].cuda(), direction)
    assert nf.ops.LAUNCHES > before
    ref = g[direction]
    n_sum = g["x"][0].numel()
    assert_close(got["z"], ref["z"], what=name + ".z." + direction, **SPLINE)
    assert_close(got["log_det"], ref["log_det"], rtol=SPLINE["rtol"], atol=SPLINE["atol"] * n_sum,
                 what=name + ".log_det." + direction)
    assert_grads_close({"x": got["gx"], **got["grads"]}, {"x": ref["gx"], **ref["grads"]}, rtol=2e-4,
                       what=name + "." + direction)


@pytest.mark.gpu
def test_bounded_spline_refuses_inputs_outside_the_unit_box():
    layer = nf.flows.neural_spline.coupling.PiecewiseRationalQuadraticCDF(shape=[3], num_bins=8).cuda()
    x = torch.rand(8, 3, device="cuda")
    layer(x)
    x[2, 1] = 1.5
    with pytest.raises(ValueError):
        layer(x)
    with pytest.raises(ValueError):
        layer.inverse(x)
