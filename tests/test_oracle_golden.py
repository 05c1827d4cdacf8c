"""Pins the oracle: every golden vector produced by the unmodified reference
(tests/golden/make_golden.py) must be reproduced by `oracle/` on CPU."""
import copy
import os

import pytest
import torch

import oracle
from cases import LAYER_CASES, MODEL_CASES
from harness import eval_layer, eval_model
from parity import assert_close, assert_grads_close

GOLD = os.path.join(os.path.dirname(__file__), "golden")
LAYERS = torch.load(os.path.join(GOLD, "layers.pt"))
MODELS = torch.load(os.path.join(GOLD, "models.pt"))


def _tol(dtype):
    return dict(rtol=1e-9, atol=1e-10) if dtype == torch.float64 else dict(rtol=2e-5, atol=2e-6)


@pytest.mark.parametrize("direction", ["inverse", "forward"])
@pytest.mark.parametrize("name", sorted(LAYER_CASES))
def test_layer_matches_reference(name, direction):
    case, gold = LAYER_CASES[name], LAYERS[name]
    dtype = case.get("dtype", torch.float32)
    torch.manual_seed(0)
    layer = case["build"](oracle).to(dtype)
    layer.load_state_dict(gold["state"])
    got = eval_layer(layer, gold["x"], direction, case.get("nograd", False))
    ref = gold[direction]
    tol = _tol(dtype)
    assert_close(got["z"], ref["z"], what=name + ".z", **tol)
    assert_close(got["log_det"], ref["log_det"], what=name + ".log_det",
                 rtol=tol["rtol"], atol=tol["atol"] * 30)
    if ref["gx"] is not None:
        assert_grads_close({"x": got["gx"], **got["grads"]}, {"x": ref["gx"], **ref["grads"]},
                           rtol=1e-8 if dtype == torch.float64 else 2e-4, what=name)


@pytest.mark.parametrize("name", sorted(MODEL_CASES))
def test_model_matches_reference(name):
    case, gold = MODEL_CASES[name], MODELS[name]
    dtype = case.get("dtype", torch.float32)
    torch.manual_seed(0)
    model = case["build"](oracle).to(dtype)
    model.load_state_dict(gold["state"])
    for call in case["calls"]:
        got = eval_model(copy.deepcopy(model), gold["x"], call, gold["y"])
        ref = gold[call]
        assert_close(got["value"], ref["value"], rtol=5e-5, atol=5e-4, what="%s.%s" % (name, call))
        assert_grads_close(got["grads"], ref["grads"], rtol=5e-4, what="%s.%s" % (name, call))
