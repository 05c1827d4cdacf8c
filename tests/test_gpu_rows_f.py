"""SURVEY.md "next" rows f2 / f4 on the GPU against golden vectors produced by the UNMODIFIED reference
(tests/golden/layers_r2.pt, `make_golden.py --r2`): 4-D spline coupling with a ConvResidualNet conditioner (the
reference's own test shape, coupling_test.py:49-50), the unconditional transform with image-shaped parameters,
checkerboard Split / Merge inside an AffineCouplingBlock, the Logit transform; plus Jitter / Scale / Shift and the
state_dict key check for the new modules."""
import os

import pytest
import torch

import vcnf_b200 as nf
from cases import LAYER_CASES_R2
from harness import eval_layer
from parity import assert_close, assert_close_tail, assert_grads_close, assert_grads_close_tail, AFFINE, SPLINE

GOLD = os.path.join(os.path.dirname(__file__), "golden", "layers_r2.pt")


@pytest.mark.parametrize("name", sorted(LAYER_CASES_R2))
def test_r2_state_dict_keys_match_reference(name):
    gold = torch.load(GOLD)[name]
    torch.manual_seed(0)
    layer = LAYER_CASES_R2[name]["build"](nf)
    layer.load_state_dict(gold["state"], strict=True)
    assert list(layer.state_dict().keys()) == list(gold["state"].keys())


@pytest.mark.gpu
@pytest.mark.parametrize("direction", ["inverse", "forward"])
@pytest.mark.parametrize("name", sorted(LAYER_CASES_R2))
def test_r2_layer_matches_reference_golden(name, direction):
    case, g = LAYER_CASES_R2[name], torch.load(GOLD)[name]
    torch.manual_seed(0)
    layer = case["build"](nf)
    layer.load_state_dict(g["state"])
    layer = layer.cuda()
    before = nf.ops.LAUNCHES
    got = eval_layer(layer, g["x"].cuda(), direction)
    assert nf.ops.LAUNCHES > before
    ref, r64 = g[direction], g[direction + "_f64"]
    n_sum = g["x"][0].numel()
    tol = SPLINE if name.startswith("prq") else AFFINE
    assert_close_tail(got["z"], ref["z"], r64["z"], what=name + ".z." + direction, **tol)
    assert_close_tail(got["log_det"], ref["log_det"], r64["log_det"], what=name + ".log_det." + direction,
                      rtol=tol["rtol"], atol=tol["atol"] * n_sum)
    if name.startswith("prq"):
        # spline gradients with the fp64 tail clause of SURVEY.md B.1 (the reference's own fp64 gradients of the same
        # case are part of the fixture).  These fixtures (10 x 2 x 4 x 4, 16 hidden channels perturbed by 0.2, K = 10
        # bins, inputs reaching 3.9 with the tails at 3) are ill-conditioned in the sampling direction: the input
        # gradient agrees to 3e-5 of its scale, the conditioner-parameter gradients -- sums dominated by a few
        # elements with a tiny spline derivative, where the kernel differentiates implicitly (dx/dphi = -f_phi / f')
        # and the reference differentiates its quadratic-root formula -- to 1.2e-4 .. 2.9e-4 (tools/diag_r2.py;
        # the reference's own fp32-vs-fp64 gap there is 2e-5).  Stated for these two cases: 5e-4.
        assert_grads_close_tail({"x": got["gx"], **got["grads"]}, {"x": ref["gx"], **ref["grads"]},
                                {"x": r64["gx"], **r64["grads"]}, rtol=5e-4 if direction == "inverse" else 2e-4,
                                what=name + "." + direction, frac=0.999, worst=4.0)
    else:
        assert_grads_close({"x": got["gx"], **got["grads"]}, {"x": ref["gx"], **ref["grads"]}, rtol=1e-4,
                           what=name + "." + direction)


@pytest.mark.gpu
def test_checkerboard_split_is_the_reference_index_map():
    """Split('checkerboard') on 2-D and 4-D inputs against the mask construction of reshape.py:30-44."""
    for shape in [(3, 6), (2, 3, 4, 6), (4, 5, 2)]:
        z = torch.randn(*shape, generator=torch.Generator().manual_seed(1))
        for mode in ("checkerboard", "checkerboard_inv"):
            n_dims = z.dim()
            cb0, cb1 = 0, 1
            for i in range(1, n_dims):
                cb0_, cb1_ = cb0, cb1
                cb0 = [cb0_ if j % 2 == 0 else cb1_ for j in range(z.size(n_dims - i))]
                cb1 = [cb1_ if j % 2 == 0 else cb0_ for j in range(z.size(n_dims - i))]
            cb = torch.tensor(cb1 if "inv" in mode else cb0)[None].repeat(len(z), *((n_dims - 1) * [1]))
            z1 = z.reshape(-1)[torch.nonzero(cb.view(-1), as_tuple=False)].view(*z.shape[:-1], -1)
            z2 = z.reshape(-1)[torch.nonzero((1 - cb).view(-1), as_tuple=False)].view(*z.shape[:-1], -1)
            layer = nf.flows.Split(mode)
            (g1, g2), ld = layer(z.cuda())
            assert ld == 0 and torch.equal(g1.cpu(), z1) and torch.equal(g2.cpu(), z2)
            back, _ = layer.inverse([g1, g2])
            assert torch.equal(back.cpu(), z)
            merged, _ = nf.flows.Merge(mode)([g1, g2])
            assert torch.equal(merged.cpu(), z)


@pytest.mark.gpu
def test_preprocessing_transforms_on_device():
    """utils.Jitter / Scale / Logit (utils/preprocessing.py:5-56) and transforms.Shift (transforms.py:42-62)."""
    x = torch.rand(7, 3, 8, 8, generator=torch.Generator().manual_seed(2))
    xg = x.cuda()
    assert torch.equal(nf.utils.Scale(255. / 256.)(xg).cpu(), x * (255. / 256.))
    torch.manual_seed(3)
    j = nf.utils.Jitter(1. / 256)(xg)
    torch.manual_seed(3)
    eps = torch.rand_like(xg)
    assert_close(j, (xg + eps * (1. / 256)), rtol=1e-6, atol=1e-7, what="jitter")
    lg = nf.utils.Logit(alpha=0.05)
    x_ = 0.05 + 0.95 * x
    assert_close(lg(xg), torch.log(x_ / (1 - x_)), rtol=1e-5, atol=1e-5, what="preproc.logit")
    assert_close(lg.inverse(lg(xg)), x, rtol=1e-5, atol=1e-5, what="preproc.logit.round_trip")
    sh = nf.transforms.Shift(-0.5)
    y = xg.clone()
    out, ld = sh(y)
    assert out.data_ptr() == y.data_ptr() and ld == 0.       # in place like the reference
    assert torch.equal(out.cpu(), x + 0.5)
    out, _ = sh.inverse(out)
    assert_close(out, x, rtol=1e-6, atol=1e-7, what="shift.round_trip")


@pytest.mark.gpu
def test_glow_multiscale_with_logit_transform_builds_and_trains():
    """example/glow.ipynb builds MultiscaleFlow(..., transform=nf.transforms.Logit()); one training step."""
    from cases import build_glow_small
    torch.manual_seed(0)
    model = build_glow_small(nf, L=2, K=2, hidden=16, in_shape=(3, 8, 8), classes=4)
    model.transform = nf.transforms.Logit(alpha=0.05)
    model = model.cuda()
    x = (torch.rand(8, 3, 8, 8) * 255 / 256 + torch.rand(8, 3, 8, 8) / 256).cuda()
    y = torch.randint(0, 4, (8,)).cuda()
    loss = model.forward_kld(x, y)
    loss.backward()
    assert torch.isfinite(loss)
    with torch.no_grad():
        z, lq = model.sample(4, y[:4])
    assert z.shape == (4, 3, 8, 8) and float(z.min()) >= -0.06 and float(z.max()) <= 1.06
