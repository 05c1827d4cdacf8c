"""optim.FlatAdam (vcnf_adam_flat, one launch on flat buffers) against torch.optim.Adam with the reference's
hyper-parameters (runcred.py:192: lr, weight_decay = 1e-6, default betas / eps)."""
import copy

import pytest
import torch

import vcnf_b200 as nf
from vcnf_b200 import ops, optim, parallel

pytestmark = pytest.mark.gpu


def _model(dtype):
    torch.manual_seed(3)
    flows = []
    for _ in range(2):
        flows += [nf.flows.CoupledRationalQuadraticSpline(5, 1, 16), nf.flows.LULinearPermute(5)]
    m = nf.NormalizingFlow(nf.distributions.DiagGaussian(5, trainable=False), flows)
    return m.to(dtype).cuda()


@pytest.mark.parametrize("dtype,wd", [(torch.float32, 0.0), (torch.float32, 1e-6), (torch.float64, 1e-6)])
def test_flat_adam_matches_torch_adam(dtype, wd):
    a = _model(dtype)
    b = copy.deepcopy(a)
    x = torch.randn(512, 5, device="cuda", dtype=dtype)
    opt_a = torch.optim.Adam(a.parameters(), lr=1e-3, weight_decay=wd)
    opt_b = optim.FlatAdam(b.parameters(), lr=1e-3, weight_decay=wd)
    assert all(p.data_ptr() >= opt_b.flat_p.data_ptr() for p in b.parameters() if p.requires_grad)
    for _ in range(6):
        opt_a.zero_grad(set_to_none=True)
        a.forward_kld(x).backward()
        opt_a.step()
        opt_b.zero_grad()
        lb = b.forward_kld(x)
        lb.backward()
        before = ops.LAUNCHES
        opt_b.step()
        assert ops.LAUNCHES == before + 1                      # one launch whatever the number of tensors
    assert float(opt_b.steps) == 6.0
    tol = 2e-5 if dtype == torch.float32 else 1e-10            # six steps of lr-sized moves, fp32 rounding
    for (n, pa), pb in zip(a.named_parameters(), b.parameters()):
        assert torch.allclose(pa, pb, rtol=tol, atol=tol), (n, float((pa - pb).abs().max()))
    assert torch.isfinite(opt_b.flat_p).all()


def test_flat_adam_single_step_is_lr_sized():
    p = torch.nn.Parameter(torch.linspace(-1, 1, 1003, device="cuda"))
    opt = optim.FlatAdam([p], lr=1e-2)
    ref = p.detach().clone()
    g = torch.randn(1003, device="cuda")
    g[::7] = 0
    p.grad = g.clone()
    opt.step()
    want = ref - 1e-2 * g / (g.abs() + 1e-8)                   # t = 1: m / (1 - b1) = g, sqrt(v / (1 - b2)) = |g|
    assert torch.allclose(p.detach(), want, rtol=1e-5, atol=1e-7)
    assert p.grad.data_ptr() == opt.bucket.flat.data_ptr()     # collect() re-bound the gradient to the bucket


def test_flat_adam_in_a_replayed_step():
    m = _model(torch.float32)
    ref = copy.deepcopy(m)
    x = torch.randn(256, 5, device="cuda")
    opt = optim.FlatAdam(m.parameters(), lr=1e-3)
    step = parallel.GraphedTrainStep(m, opt, opt.bucket, x, world=1, warmup=3)
    assert step.graph is not None, step.capture_error
    for _ in range(5):
        loss = step(x)
    torch.cuda.synchronize()
    assert float(opt.steps) == 5                               # replays only: warm-ups and the capture do not train
    opt_r = torch.optim.Adam(ref.parameters(), lr=1e-3)
    for _ in range(int(opt.steps)):
        opt_r.zero_grad(set_to_none=True)
        ref.forward_kld(x).backward()
        opt_r.step()
    for pa, pb in zip(ref.parameters(), m.parameters()):
        assert torch.allclose(pa, pb, rtol=5e-5, atol=5e-5), float((pa - pb).abs().max())
