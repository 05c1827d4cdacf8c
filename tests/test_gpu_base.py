"""Base-distribution kernels (vcnf_gauss_fwd / _bwd, vcnf_gmm_logprob / _bwd) against the oracle's
distributions (oracle/distributions.py restates normflow/distributions/base.py:609-652, 715-775, 957-1034) on the
CPU: values 1e-5 relative (fp32) / 1e-10 (fp64), gradients 1e-4 / 1e-9 of the tensor scale."""
import copy

import pytest
import torch

import oracle
import vcnf_b200 as nf
from vcnf_b200 import ops
from parity import assert_close, assert_grads_close

pytestmark = pytest.mark.gpu


def _g(seed):
    return torch.Generator().manual_seed(seed)


def _perturb(m, seed=3, scale=0.3):
    with torch.no_grad():
        for _, p in sorted(m.named_parameters()):
            p.add_(scale * torch.randn(p.shape, generator=_g(seed)).to(p.dtype))
    return m


def _grads(model, loss, x=None):
    names = [n for n, p in model.named_parameters() if p.requires_grad]
    ps = [p for _, p in model.named_parameters() if p.requires_grad]
    gs = torch.autograd.grad(loss, ps + ([x] if x is not None else []), allow_unused=True)
    out = dict(zip(names, gs[:len(ps)]))
    if x is not None:
        out["x"] = gs[-1]
    return out


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
@pytest.mark.parametrize("shape,B", [(29, 4099), ((6, 16, 16), 37), (2, 1), ((48, 4, 4), 130)])
def test_diag_gaussian_log_prob(shape, B, dtype):
    ref = _perturb(oracle.distributions.DiagGaussian(shape)).to(dtype)
    prod = nf.distributions.DiagGaussian(shape).to(dtype)
    prod.load_state_dict(ref.state_dict())
    prod = prod.cuda()
    shp = (shape,) if isinstance(shape, int) else tuple(shape)
    x = torch.randn((B,) + shp, generator=_g(1)).to(dtype)
    w = torch.randn(B, generator=_g(2)).to(dtype)
    xr, xg = x.clone().requires_grad_(True), x.cuda().requires_grad_(True)
    lr, lg = ref.log_prob(xr), prod.log_prob(xg)
    tol = dict(rtol=1e-5, atol=1e-5) if dtype == torch.float32 else dict(rtol=1e-10, atol=1e-10)
    assert_close(lg, lr, what="diag.log_prob", **tol)
    gr, gg = _grads(ref, (lr * w).sum(), xr), _grads(prod, (lg * w.cuda()).sum(), xg)
    assert_grads_close(gg, gr, rtol=1e-4 if dtype == torch.float32 else 1e-9, what="diag.log_prob")
    # temperature annealing (base.py:636-640)
    ref.temperature = prod.temperature = 0.7
    assert_close(prod.log_prob(x.cuda()), ref.log_prob(x), what="diag.log_prob(T)", **tol)


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_class_cond_gaussian_log_prob_and_sample(dtype):
    shape, classes, B = (6, 8, 8), 10, 53
    ref = _perturb(oracle.distributions.ClassCondDiagGaussian(shape, classes)).to(dtype)
    prod = nf.distributions.ClassCondDiagGaussian(shape, classes).to(dtype)
    prod.load_state_dict(ref.state_dict())
    prod = prod.cuda()
    x = torch.randn((B,) + shape, generator=_g(4)).to(dtype)
    y = torch.randint(0, classes, (B,), generator=_g(5))
    w = torch.randn(B, generator=_g(6)).to(dtype)
    xr, xg = x.clone().requires_grad_(True), x.cuda().requires_grad_(True)
    lr, lg = ref.log_prob(xr, y), prod.log_prob(xg, y.cuda())
    tol = dict(rtol=1e-5, atol=1e-5) if dtype == torch.float32 else dict(rtol=1e-10, atol=1e-10)
    assert_close(lg, lr, what="classcond.log_prob", **tol)
    gr, gg = _grads(ref, (lr * w).sum(), xr), _grads(prod, (lg * w.cuda()).sum(), xg)
    assert_grads_close(gg, gr, rtol=1e-4 if dtype == torch.float32 else 1e-9, what="classcond.log_prob")
    # reparametrised draw from given standard-normal noise: z = loc_y + exp(log_scale_y) eps (base.py:737-752)
    eps = torch.randn((B,) + shape, generator=_g(7)).to(dtype)
    z, lp = ops.gauss_sample(eps.cuda(), prod.loc, prod.log_scale, y=y.cuda(), classes=classes)
    onehot = torch.nn.functional.one_hot(y, classes).to(dtype).t()
    loc = (ref.loc @ onehot).permute(3, 0, 1, 2)
    ls = (ref.log_scale @ onehot).permute(3, 0, 1, 2)
    zr = loc + torch.exp(ls) * eps
    assert_close(z, zr, what="classcond.sample.z", **tol)
    assert_close(lp, ref.log_prob(zr, y), what="classcond.sample.log_p", **tol)
    gz, gl = torch.randn(z.shape, generator=_g(8)).to(dtype), torch.randn(B, generator=_g(9)).to(dtype)
    lpr = -0.5 * ref.d * torch.log(torch.tensor(2 * torch.pi, dtype=dtype)) - (ls + 0.5 * eps ** 2).sum((1, 2, 3))
    gr = _grads(ref, (zr * gz).sum() + (lpr * gl).sum())
    gg = _grads(prod, (z * gz.cuda()).sum() + (lp * gl.cuda()).sum())
    assert_grads_close(gg, gr, rtol=1e-4 if dtype == torch.float32 else 1e-9, what="classcond.sample")


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
@pytest.mark.parametrize("K,d,B", [(100, 15, 2048), (4, 6, 33), (7, 2, 1)])
def test_gaussian_mixture_log_prob(K, d, B, dtype):
    ref = _perturb(oracle.distributions.GaussianMixture(n_modes=K, dim=d), scale=0.5).to(dtype)
    prod = nf.distributions.GaussianMixture(n_modes=K, dim=d).to(dtype)
    prod.load_state_dict(ref.state_dict())
    prod = prod.cuda()
    x = (1.5 * torch.randn(B, d, generator=_g(10))).to(dtype)
    w = torch.randn(B, generator=_g(11)).to(dtype)
    xr, xg = x.clone().requires_grad_(True), x.cuda().requires_grad_(True)
    lr, lg = ref.log_prob(xr), prod.log_prob(xg)
    tol = dict(rtol=1e-5, atol=1e-5) if dtype == torch.float32 else dict(rtol=1e-10, atol=1e-10)
    assert_close(lg, lr, what="gmm.log_prob", **tol)
    gr, gg = _grads(ref, (lr * w).sum(), xr), _grads(prod, (lg * w.cuda()).sum(), xg)
    assert_grads_close(gg, gr, rtol=1e-4 if dtype == torch.float32 else 1e-9, what="gmm.log_prob")


def test_diag_gaussian_draw_is_consistent():
    """DiagGaussian.forward on the device: log_p of the draw equals log_prob of the returned points, and the
    reparametrised gradient reaches loc / log_scale (cfg 1 trains them through reverse_kld)."""
    q0 = _perturb(nf.distributions.DiagGaussian(2)).cuda()
    z, lp = q0(4096)
    assert_close(lp, q0.log_prob(z.detach()), rtol=1e-5, atol=1e-5, what="diag.draw")
    (z.sum() + lp.sum()).backward()
    assert float((q0.loc.grad - 4096).abs().max()) < 1e-2
    assert q0.log_scale.grad is not None and torch.isfinite(q0.log_scale.grad).all()


def test_class_cond_flow_container():
    """nf.ClassCondFlow (core.py:200-268): the loops around a class-conditional base, against the same layers applied by
    hand; sample() and log_prob() are consistent with each other."""
    torch.manual_seed(2)
    flows = [nf.flows.MaskedAffineFlow(torch.tensor([1., 0., 1., 0.]), nf.nets.MLP([4, 8, 4]), nf.nets.MLP([4, 8, 4])),
             nf.flows.Permute(4, mode="swap")]
    model = nf.ClassCondFlow(nf.distributions.ClassCondDiagGaussian(4, 3), flows).cuda()
    x = torch.randn(257, 4, device="cuda")
    y = torch.randint(0, 3, (257,), device="cuda")
    z, ld = x, torch.zeros(257, device="cuda")
    for f in reversed(model.flows):
        z, d = f.inverse(z)
        ld = ld + d
    want = ld + model.q0.log_prob(z, y)
    got = model.log_prob(x, y)
    assert torch.allclose(got, want, rtol=1e-6, atol=1e-6)
    assert abs(float(model.forward_kld(x, y)) + float(want.mean())) < 1e-5
    with torch.no_grad():
        xs, lq = model.sample(64, y[:64])
        assert xs.shape == (64, 4)
        assert torch.allclose(model.log_prob(xs, y[:64]), lq, rtol=1e-4, atol=1e-4)
