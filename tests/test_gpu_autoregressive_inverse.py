"""Sampling direction of the autoregressive layers (affine/autoregressive.py:24-36 of the reference: D conditioner
passes).  Without a gradient the package evaluates only the final-layer rows and the one column a pass can fix;
these tests pin that route to the reference's full loop (`full_inverse = True`)."""
import pytest
import torch

import vcnf_b200 as nf
from vcnf_b200 import ops

pytestmark = pytest.mark.gpu


def _randomize(m, scale):
    g = torch.Generator().manual_seed(5)
    with torch.no_grad():
        for p in m.parameters():
            p.add_(torch.randn(p.shape, generator=g).to(p) * scale)


def _both(layer, z, fused=False):
    with torch.no_grad():
        layer.full_inverse = True
        x_ref, ld_ref = layer.inverse(z)
        layer.full_inverse = False
        layer.fused_inverse = fused          # False: the sliced passes (bit-identical); True: vcnf_made_inverse
        before = ops.LAUNCHES
        x, ld = layer.inverse(z)
        launches = ops.LAUNCHES - before
        layer.fused_inverse = True
    return x_ref, ld_ref, x, ld, launches


@pytest.mark.parametrize("kind,D,H,permute,bins,rows", [
    ("spline", 8, 32, False, 8, 1003), ("spline", 8, 32, True, 8, 1003), ("spline", 64, 256, False, 8, 1003),
    ("spline", 16, 64, True, 10, 1003), ("maf", 16, 64, False, 0, 1003), ("maf", 64, 256, False, 0, 1003),
    # several 16-row tiles per CTA (148 CTAs x 16 rows per wave): the weight hand-over between tiles and between the
    # two sample groups of a CTA
    ("maf", 64, 256, False, 0, 20011), ("spline", 64, 256, False, 8, 20011), ("maf", 16, 64, False, 0, 50000)])
def test_fused_inverse_kernel_equals_the_full_loop(kind, D, H, permute, bins, rows):
    """vcnf_made_inverse (all D passes in one launch, one MADE evaluation in total) against the reference's D-pass
    loop on the same weights: same dot products in a different summation order -> 1e-5 of the scale for the values,
    1e-4 for the log-det (a sum of D per-feature terms); ragged batch (not a multiple of the 32-row tile)."""
    torch.manual_seed(0)
    if kind == "spline":
        flow = nf.flows.AutoregressiveRationalQuadraticSpline(D, 2, H, num_bins=bins, permute_mask=permute).cuda()
        layer = flow.mprqat
    else:
        layer = nf.flows.MaskedAffineAutoregressive(D, H).cuda()
    _randomize(layer, 0.05)
    z = 1.5 * torch.randn(rows, D, device="cuda")
    x_ref, ld_ref, x, ld, launches = _both(layer, z, fused=True)
    if D % 16 == 0:
        assert launches == 1, launches       # the single-launch kernel splits D and H over 16 warps
    assert torch.isfinite(x).all() and torch.isfinite(ld).all()
    assert float((x - x_ref).abs().max()) <= 1e-5 * max(1.0, float(x_ref.abs().max())), float((x - x_ref).abs().max())
    assert float((ld - ld_ref).abs().max()) <= 1e-4 * max(1.0, float(ld_ref.abs().max())), float((ld - ld_ref).abs().max())
    with torch.no_grad():                                   # and it inverts the density direction
        z_back, ld_f = layer.forward(x)
    assert float((z_back - z).abs().max()) < 2e-3
    assert float((ld_f + ld).abs().max()) < 2e-3


@pytest.mark.parametrize("D,H,permute", [(8, 32, False), (8, 32, True), (64, 256, False)])
def test_spline_autoregressive_sliced_inverse_equals_full_loop(D, H, permute):
    torch.manual_seed(0)
    flow = nf.flows.AutoregressiveRationalQuadraticSpline(D, 2, H, permute_mask=permute).cuda()
    _randomize(flow, 0.05)
    layer = flow.mprqat
    z = 1.5 * torch.randn(1000, D, device="cuda")
    x_ref, ld_ref, x, ld, _ = _both(layer, z)
    assert torch.isfinite(x_ref).all()
    assert torch.equal(x, x_ref), float((x - x_ref).abs().max())
    assert torch.equal(ld, ld_ref), float((ld - ld_ref).abs().max())
    with torch.no_grad():                                   # and it inverts the density direction
        z_back, ld_f = layer.forward(x)
    assert float((z_back - z).abs().max()) < 2e-3
    assert float((ld_f + ld).abs().max()) < 2e-3


def test_affine_autoregressive_sliced_inverse_equals_full_loop():
    torch.manual_seed(1)
    layer = nf.flows.MaskedAffineAutoregressive(16, 64).cuda()
    _randomize(layer, 0.05)
    z = torch.randn(777, 16, device="cuda")
    x_ref, ld_ref, x, ld, _ = _both(layer, z)
    assert torch.equal(x, x_ref), float((x - x_ref).abs().max())
    assert torch.equal(ld, ld_ref)


def test_sliced_inverse_fp64_and_gradient_route():
    torch.manual_seed(2)
    flow = nf.flows.AutoregressiveRationalQuadraticSpline(6, 1, 16).double().cuda()
    _randomize(flow, 0.05)
    layer = flow.mprqat
    z = torch.randn(300, 6, device="cuda", dtype=torch.float64)
    x_ref, ld_ref, x, ld, _ = _both(layer, z)
    assert torch.allclose(x, x_ref, rtol=1e-10, atol=1e-10) and torch.allclose(ld, ld_ref, rtol=1e-10, atol=1e-10)
    # with a gradient required the reference's loop runs and stays differentiable
    zg = z.clone().requires_grad_(True)
    xg, ldg = layer.inverse(zg)
    (xg.sum() + ldg.sum()).backward()
    assert zg.grad is not None and torch.isfinite(zg.grad).all()
    assert torch.allclose(xg.detach(), x_ref, rtol=1e-10, atol=1e-10)
