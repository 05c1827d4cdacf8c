"""Tolerance helpers shared by the CPU (oracle vs golden) and GPU (CUDA vs oracle) tests.

Stated fp32 tolerances (BASELINE.json north_star):
  affine family : |a-b| <= 1e-6 + 1e-5*|b|
  splines       : |a-b| <= 1e-5 + 1e-4*|b|   (log-dets: atol scaled by the number of summed elements)
"""
import torch

AFFINE = dict(rtol=1e-5, atol=1e-6)
SPLINE = dict(rtol=1e-4, atol=1e-5)
F64 = dict(rtol=1e-9, atol=1e-11)


def excess(a, b, rtol, atol):
    """max over elements of |a-b| / (atol + rtol*|b|); <= 1 passes."""
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    if a.numel() == 0:
        return 0.0
    nan_mismatch = (torch.isnan(a) != torch.isnan(b)).any().item()
    if nan_mismatch:
        return float("inf")
    ok = ~torch.isnan(b)
    return float(((a - b).abs()[ok] / (atol + rtol * b.abs()[ok])).max()) if ok.any() else 0.0


def assert_close(a, b, rtol, atol, what=""):
    assert tuple(a.shape) == tuple(b.shape), (what, a.shape, b.shape)
    e = excess(a, b, rtol, atol)
    assert e <= 1.0, "%s: error is %.3g x tolerance (rtol=%g atol=%g, max|ref|=%.3g)" % (
        what, e, rtol, atol, float(b.abs().max()) if b.numel() else 0.0)


def assert_grads_close(got, ref, rtol, what=""):
    """Gradient tensors: atol scaled by the tensor's max-abs (SURVEY.md B.1 last bullet)."""
    for name, r in ref.items():
        g = got.get(name)
        if r is None:
            assert g is None or float(g.abs().max()) == 0.0, (what, name)
            continue
        assert g is not None, (what, name, "missing gradient")
        scale = float(r.abs().max()) if r.numel() else 0.0
        assert_close(g, r, rtol=rtol, atol=rtol * max(scale, 1e-30), what="%s grad[%s]" % (what, name))


def assert_close_tail(a, ref32, ref64, rtol, atol, what="", frac=0.999):
    """fp32 spline outputs: >= `frac` of the elements inside the stated tolerance of
    the fp32 reference; the ill-conditioned remainder must be no further from the fp64 reference
    than twice the fp32 reference's own worst error against it (+ 10 atol)."""
    assert tuple(a.shape) == tuple(ref32.shape), (what, a.shape, ref32.shape)
    a_, r32, r64 = a.detach().double().cpu(), ref32.detach().double().cpu(), ref64.detach().double().cpu()
    assert not (torch.isnan(a_) != torch.isnan(r32)).any(), what + ": NaN pattern differs"
    ok = ~torch.isnan(r32)
    if not ok.any():
        return
    err = (a_ - r32).abs()[ok]
    inside = err <= atol + rtol * r32.abs()[ok]
    share = float(inside.double().mean())
    e64 = float((a_ - r64).abs()[ok].max())
    ref_e = float((r32 - r64).abs()[ok].max())
    assert share >= frac, "%s: only %.4f%% of elements inside rtol=%g atol=%g (worst %.3g x tol)" % (
        what, 100 * share, rtol, atol, float((err / (atol + rtol * r32.abs()[ok])).max()))
    assert e64 <= 2 * ref_e + 10 * atol, "%s: max error vs fp64 reference %.3g > 2 x %.3g + 10 atol" % (what, e64, ref_e)
