"""CPU-side checks of the drop-in boundary: the C-ABI library loads and exports every symbol
include/vcnf_b200.h declares, the product modules load the reference's state_dicts key for key,
and the product refuses to compute without CUDA (no CPU fallback)."""
import ctypes
import os
import re

import pytest
import torch

import vcnf_b200 as nf
from vcnf_b200 import _lib
from cases import LAYER_CASES, MODEL_CASES

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")


@pytest.fixture(scope="module")
def built():
    from vcnf_b200 import build
    return build.build()


def _declared():
    src = open(os.path.join(ROOT, "include", "vcnf_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(vcnf_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol(built):
    handle = ctypes.CDLL(built)
    names = _declared()
    assert len(names) >= 15
    for n in names:
        assert hasattr(handle, n), n
    assert sorted(_lib.EXPORTED) == names          # the ctypes table covers the whole header
    assert handle.vcnf_version() == 100


def test_bad_arguments_are_rejected_without_a_gpu(built):
    lib = _lib.lib()
    assert lib.vcnf_gather_cols(None, 0, None, 0, 4, 2, None, 0, None) == -1
    assert lib.vcnf_small_linear(None, None, None, None, None, None, 1, 4, 1, 0, 0, None) == -1
    assert lib.vcnf_squeeze(None, None, 1, 1, 1, 1, 0, 0, None) == -1


@pytest.mark.parametrize("name", sorted(LAYER_CASES))
def test_layer_state_dict_keys_match_reference(name):
    gold = torch.load(os.path.join(GOLD, "layers.pt"))[name]
    torch.manual_seed(0)
    layer = LAYER_CASES[name]["build"](nf)
    layer.load_state_dict(gold["state"], strict=True)
    assert list(layer.state_dict().keys()) == list(gold["state"].keys())


@pytest.mark.parametrize("name", sorted(MODEL_CASES))
def test_model_state_dict_keys_match_reference(name):
    gold = torch.load(os.path.join(GOLD, "models.pt"))[name]
    torch.manual_seed(0)
    model = MODEL_CASES[name]["build"](nf)
    model.load_state_dict(gold["state"], strict=True)
    assert list(model.state_dict().keys()) == list(gold["state"].keys())


def test_no_cpu_fallback():
    layer = nf.flows.CoupledRationalQuadraticSpline(4, 1, 8)
    with pytest.raises(RuntimeError, match="CUDA"):
        layer.inverse(torch.randn(3, 4))
    with pytest.raises(RuntimeError, match="CUDA"):
        nf.flows.AffineCouplingBlock(nf.nets.MLP([1, 4, 2]))(torch.randn(3, 2))


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "vcnf_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(".py"):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f


def test_fused_net_shape_queries_run_without_a_gpu():
    """vcnf_resnet_spline_pack_bytes is pure host code: supported shapes give a size, unsupported ones VCNF_E_UNSUPPORTED."""
    from vcnf_b200 import _lib
    assert _lib.query("vcnf_resnet_spline_pack_bytes", 15, 128, 2, 14, 23) == (1 + 8 * 7) * 16384 + 8 * 128 * 4
    assert _lib.query("vcnf_resnet_spline_pack_bytes", 15, 64, 2, 14, 23) == -2       # hidden must be 128
    assert _lib.query("vcnf_resnet_spline_pack_bytes", 40, 128, 2, 14, 23) == -2      # more than 32 net inputs
    assert _lib.query("vcnf_resnet_spline_pack_bytes", 15, 128, 2, 32, 25) == -2      # 32 features x 32 columns > 4 slices


def test_fused_route_is_a_cuda_only_choice():
    """Host logic of the fused conditioner + spline route: never taken for host tensors (the layer-by-layer ops then raise
    the usual 'no CPU path' error), its specification is built from the layer's own column lists and tail modes."""
    import torch
    import vcnf_b200 as nf
    from vcnf_b200 import ops
    layer = nf.flows.CoupledRationalQuadraticSpline(29, 2, 128)
    c = layer.prqct
    x = torch.randn(4, 29)
    assert c._fused_route(x, None) is None
    spec = ops.FusedNetSpec(c.identity_features.tolist(), c.transform_features.tolist(), [ops.TAIL_LINEAR] * 14, [3.0] * 14,
                            8, 23, 1.0 / 128 ** 0.5, False, 128, 2, [ops.TAIL_LINEAR] * 15, [3.0] * 15, 23)
    assert spec.supported and spec.n_in == 15 and spec.Tc == 14 and spec.Ts == 15
    assert list(spec.id_cols) == c.identity_features.tolist() and list(spec.tr_cols) == c.transform_features.tolist()
    assert not ops.FusedNetSpec([0], [1], [ops.TAIL_LINEAR], [3.0], 8, 23, 1.0, False, 64, 2).supported      # hidden 64
    assert not ops.FusedNetSpec([0], [1], [ops.TAIL_LINEAR], [3.0], 10, 29, 1.0, False, 128, 2).supported    # 10 bins
