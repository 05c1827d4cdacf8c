"""Tolerance helpers shared by the CPU (oracle vs golden) and GPU (CUDA vs oracle) tests.

Stated fp32 tolerances (BASELINE.json north_star):
  affine family : |a-b| <= 1e-6 + 1e-5*|b|
  splines       : |a-b| <= 1e-5 + 1e-4*|b|   (log-dets: atol scaled by the number of summed elements)
"""
import torch

AFFINE = dict(rtol=1e-5, atol=1e-6)
SPLINE = dict(rtol=1e-4, atol=1e-5)
F64 = dict(rtol=1e-9, atol=1e-11)


# observed maxima per check (worst error / tolerance), written out by tests/conftest.py at session end
OBSERVED = {}


def record(what, value):
    if what:
        OBSERVED[what] = max(float(value), OBSERVED.get(what, 0.0))


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
    record(what, e)
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


def assert_close_tail(a, ref32, ref64, rtol, atol, what="", frac=0.9999, worst=2.0):
    """fp32 outputs: >= `frac` of the elements inside the stated tolerance of the
    reference (its fp32 or its fp64 evaluation); the ill-conditioned remainder must be no further from the
    fp64 reference than `worst` (2) x the fp32 reference's own worst error against it (+ 10 atol).  (Rounding errors on
    such elements are amplified by the same huge factor for any fp32 evaluation order; which order gets
    the luckier draw on a given element is chance -- golden case crqs_linear: the reference itself is off
    by 0.42 in one per-sample log-det.)"""
    assert tuple(a.shape) == tuple(ref32.shape), (what, a.shape, ref32.shape)
    a_, r32, r64 = a.detach().double().cpu(), ref32.detach().double().cpu(), ref64.detach().double().cpu()
    assert not (torch.isnan(a_) != torch.isnan(r32)).any(), what + ": NaN pattern differs"
    ok = ~torch.isnan(r32)
    if not ok.any():
        return
    err = (a_ - r32).abs()[ok]
    tol = atol + rtol * r32.abs()[ok]
    err64 = (a_ - r64).abs()[ok]
    # an element agrees when it is inside the stated tolerance of the reference evaluated in fp32 OR in fp64
    # (the latter is the stronger statement: the reference's fp32 value is itself only an approximation of it)
    near = (err <= tol) | (err64 <= atol + rtol * r64.abs()[ok])
    # elements on which the reference's OWN fp32 rounding already uses a quarter of the tolerance are
    # ill-conditioned: a different (equally valid) fp32 evaluation order cannot be held to `tol` there;
    # they are covered by the worst-error clause below instead
    excused = (r32 - r64).abs()[ok] > 0.25 * tol
    share = float((near | excused).double().mean())
    record(what + " (worst err / tol, tail clause)", float((err / tol).max()))
    record(what + " (share outside tol, %)", 100.0 * (1.0 - float(near.double().mean())))
    e64 = float(err64[~near].max()) if (~near).any() else 0.0
    ref_e = float((r32 - r64).abs()[ok].max())
    assert share >= frac, "%s: only %.4f%% of elements inside rtol=%g atol=%g (worst %.3g x tol)" % (
        what, 100 * share, rtol, atol, float((err / tol).max()))
    record(what + " (worst err vs fp64 / reference's own fp32-vs-fp64 worst)", e64 / ref_e if ref_e > 0 else 0.0)
    assert e64 <= worst * ref_e + 10 * atol, "%s: max error vs fp64 reference %.3g > %g x %.3g + 10 atol" % (what, e64, worst, ref_e)


def assert_grads_close_tail(got, ref32, ref64, rtol, what="", frac=0.9999, worst=2.0):
    """Gradients of an fp32 spline layer: an element passes when it is inside the stated tolerance
    of the fp32 reference, OR no further from the fp64 reference than twice the fp32 reference's own
    error there (+ the same tolerance).  The second clause exists because the reference's fp32
    gradient is itself wrong (even in sign) on ill-conditioned samples -- measured on the golden
    case crqs_linear/forward row 45: reference fp32 -227.5, reference fp64 +135.3."""
    for name, r32 in ref32.items():
        g = got.get(name)
        if r32 is None:
            assert g is None or float(g.abs().max()) == 0.0, (what, name)
            continue
        assert g is not None, (what, name, "missing gradient")
        g_, r32_, r64_ = g.detach().double().cpu(), r32.detach().double().cpu(), ref64[name].detach().double().cpu()
        scale = float(r32_.abs().max()) if r32_.numel() else 0.0
        tol = rtol * max(scale, 1e-30) + rtol * r32_.abs()
        ok = ((g_ - r32_).abs() <= tol) | ((g_ - r64_).abs() <= worst * (r32_ - r64_).abs() + tol)
        # population clause (rounding errors are independent per element, amplified by 1/f' on the
        # ill-conditioned ones): >= 99.9 % inside, and the worst error against fp64 no larger than
        # twice the fp32 reference's own worst error against fp64
        share = float(ok.double().mean()) if ok.numel() else 1.0
        if ok.numel():
            record("%s grad[%s] (worst err / (rtol*max|ref|), tail clause)" % (what, name),
                   float(((g_ - r32_).abs() / tol).max()))
        e64 = float((g_ - r64_).abs().max()) if ok.numel() else 0.0
        ref_e = float((r32_ - r64_).abs().max()) if ok.numel() else 0.0
        record("%s grad[%s] (share outside both clauses, %%)" % (what, name), 100.0 * (1.0 - share))
        assert share >= frac and e64 <= worst * ref_e + float(tol.max()), (
            "%s grad[%s]: %d of %d elements outside both clauses; worst vs fp64 %.3g (reference fp32: %.3g)" % (
                what, name, int((~ok).sum()), ok.numel(), e64, ref_e))
