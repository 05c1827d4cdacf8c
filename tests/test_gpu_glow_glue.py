"""The small kernels around Glow's convolutions (cfg 5): LU-parametrised 1x1 weight (mixing.py:83-102), the per-pixel
map's gradients on image tensors (mixing.py:101-128), the fused ConvNet2d weight packs, vectorised bf16 patch matrix /
column sums, Adamax on flat buffers (example/glow.ipynb:88) -- each against the library ops it replaces."""
import copy

import pytest
import torch
import torch.nn.functional as F

import vcnf_b200 as nf
from vcnf_b200 import ops, optim, _lib
from vcnf_b200._lib import ptr, stream

pytestmark = pytest.mark.gpu


def _g(seed):
    return torch.Generator().manual_seed(seed)


@pytest.mark.parametrize("C,dtype,tol", [(12, torch.float32, 2e-6), (48, torch.float32, 5e-6), (24, torch.float64, 1e-13),
                                         (64, torch.float64, 1e-13), (1, torch.float32, 1e-6)])
def test_glow_w_matches_the_library_ops(C, dtype, tol):
    torch.manual_seed(C)
    layer = nf.flows.Invertible1x1Conv(C, use_lu=True).to(dtype).cuda()
    with torch.no_grad():
        layer.L.add_(0.1 * torch.randn_like(layer.L))
        layer.U.add_(0.1 * torch.randn_like(layer.U))
        layer.log_S.add_(0.1 * torch.randn_like(layer.log_S))
    ref = copy.deepcopy(layer)
    gW = torch.randn(C, C, generator=_g(1)).to(dtype).cuda()
    Wr = ref._assemble_W()
    ldr = 7.0 * torch.sum(ref.log_S)
    ((Wr * gW).sum() + 1.3 * ldr).backward()
    W, ld = ops.glow_w(layer.L, layer.U, layer.log_S, layer.P, layer.sign_S, 7.0)
    ((W * gW).sum() + 1.3 * ld).backward()
    scale = float(Wr.abs().max())
    assert float((W - Wr).abs().max()) <= tol * scale
    assert abs(float(ld - ldr)) <= tol * max(1.0, abs(float(ldr)))
    for name in ("L", "U", "log_S"):
        a, b = getattr(layer, name).grad, getattr(ref, name).grad
        assert float((a - b).abs().max()) <= 10 * tol * max(1.0, float(b.abs().max())), name


@pytest.mark.parametrize("B,C,H,W,dtype", [(5, 12, 16, 16, torch.float32), (3, 48, 4, 4, torch.float32),
                                           (2, 24, 8, 8, torch.float64), (1, 7, 3, 5, torch.float32)])
def test_invertible_1x1_conv_training_direction(B, C, H, W, dtype):
    """Invertible1x1Conv.inverse (the density direction) with all gradients against conv2d + autograd in fp64."""
    torch.manual_seed(B + C)
    layer = nf.flows.Invertible1x1Conv(C, use_lu=True).to(dtype).cuda()
    with torch.no_grad():
        layer.L.add_(0.05 * torch.randn_like(layer.L))
        layer.U.add_(0.05 * torch.randn_like(layer.U))
    ref = copy.deepcopy(layer).double().cpu()
    z = torch.randn(B, C, H, W, generator=_g(2)).to(dtype)
    gz = torch.randn(B, C, H, W, generator=_g(3)).to(dtype)
    zc = z.cuda().requires_grad_(True)
    y, ld = layer.inverse(zc)
    ((y * gz.cuda()).sum() + 0.7 * ld).backward()
    zr = z.double().requires_grad_(True)
    Wr = ref._assemble_W()
    yr = F.conv2d(zr, Wr.view(C, C, 1, 1))
    ldr = torch.sum(ref.log_S) * H * W
    ((yr * gz.double()).sum() + 0.7 * ldr).backward()
    tol = 2e-5 if dtype == torch.float32 else 1e-11
    rel = lambda a, b: float((a.detach().double().cpu() - b).abs().max() / (b.abs().max() + 1e-30))   # noqa: E731
    assert rel(y, yr.detach()) <= tol and abs(float(ld) - float(ldr)) <= tol * max(1.0, abs(float(ldr)))
    assert rel(zc.grad, zr.grad) <= tol
    for name in ("L", "U", "log_S"):
        assert rel(getattr(layer, name).grad, getattr(ref, name).grad) <= 5 * tol, name


def test_convnet3_pack_layouts():
    C0, Hd, C3 = 6, 64, 12
    w1 = torch.randn(Hd, C0, 3, 3, generator=_g(4)).cuda()
    w2 = torch.randn(Hd, Hd, 1, 1, generator=_g(5)).cuda()
    w3 = torch.randn(C3, Hd, 3, 3, generator=_g(6)).cuda()
    b3 = torch.randn(C3, generator=_g(7)).cuda()
    K1p, Np3, K3p, Np1 = 64, 16, 128, 16
    n = _lib.query("vcnf_convnet3_pack_elems", C0, Hd, C3, 1)
    sizes = [Hd * K1p, Hd * Hd, 9 * Np3 * Hd, Hd * K3p, Hd * Hd, 9 * Np1 * Hd]
    assert n == sum(sizes)
    pack = torch.empty(n, dtype=torch.bfloat16, device="cuda")
    b3p = torch.empty(Np3, device="cuda")
    ops._launch("vcnf_convnet3_pack", ptr(w1), ptr(w2), ptr(w3), ptr(b3), C0, Hd, C3, 1, ptr(pack), ptr(b3p), stream())
    seg = torch.split(pack, sizes)
    bf, pad = torch.bfloat16, F.pad
    want = [pad(w1.permute(0, 2, 3, 1).reshape(Hd, 9 * C0), (0, K1p - 9 * C0)),
            w2.reshape(Hd, Hd),
            pad(w3.permute(2, 3, 0, 1).reshape(9, C3, Hd), (0, 0, 0, Np3 - C3)),
            pad(w3.permute(1, 2, 3, 0).reshape(Hd, 9 * C3), (0, K3p - 9 * C3)),
            w2.reshape(Hd, Hd).t(),
            pad(w1.flip(2, 3).permute(2, 3, 1, 0).reshape(9, C0, Hd), (0, 0, 0, Np1 - C0))]
    for i, (s, w) in enumerate(zip(seg, want)):
        assert torch.equal(s, w.to(bf).reshape(-1)), i
    assert torch.equal(b3p, pad(b3, (0, Np3 - C3)))


@pytest.mark.parametrize("B,C,H,W,flip", [(3, 6, 16, 16, False), (2, 12, 8, 8, True), (5, 24, 4, 4, True)])
def test_im2col_bf16_vector_path(B, C, H, W, flip):
    x = torch.randn(B, C, H, W, generator=_g(8))
    u = F.unfold(x.flip(2, 3) if flip else x, 3, padding=1)
    if flip:
        u = u.view(B, C * 9, H, W).flip(2, 3).reshape(B, C * 9, H * W)
    ref = u.view(B, C, 9, H * W).permute(0, 3, 2, 1).reshape(B * H * W, 9 * C)
    Kp = (9 * C + 63) // 64 * 64
    col = ops.im2col_raw(x.cuda(), "nchw", B, C, H, W, 3, Kp, flip=flip, bf16=True)
    assert torch.equal(col[:, :9 * C].cpu(), ref.to(torch.bfloat16))
    assert float(col[:, 9 * C:].float().abs().max()) == 0.0


@pytest.mark.parametrize("P,C", [(4099, 256), (777, 64), (300, 24), (5, 8)])
def test_colsum_bf16_vector_path(P, C):
    x = torch.randn(P, C, generator=_g(9)).to(torch.bfloat16).cuda()
    got = ops.colsum(x, 0, P, C, C)
    want = x.double().sum(0)
    assert float((got.double() - want).abs().max()) <= 1e-4 * max(1.0, float(want.abs().max()))
    assert torch.equal(got, ops.colsum(x, 0, P, C, C))                 # fixed order: bit-reproducible


@pytest.mark.parametrize("dtype,wd", [(torch.float32, 0.0), (torch.float32, 1e-5), (torch.float64, 1e-5)])
def test_flat_adamax_matches_torch_adamax(dtype, wd):
    torch.manual_seed(3)
    flows = []
    for _ in range(2):
        flows += [nf.flows.CoupledRationalQuadraticSpline(5, 1, 16), nf.flows.LULinearPermute(5)]
    a = nf.NormalizingFlow(nf.distributions.DiagGaussian(5, trainable=False), flows).to(dtype).cuda()
    b = copy.deepcopy(a)
    x = torch.randn(512, 5, device="cuda", dtype=dtype)
    opt_a = torch.optim.Adamax(a.parameters(), lr=1e-3, weight_decay=wd)
    opt_b = optim.FlatAdamax(b.parameters(), lr=1e-3, weight_decay=wd)
    for _ in range(6):
        opt_a.zero_grad(set_to_none=True)
        a.forward_kld(x).backward()
        opt_a.step()
        opt_b.zero_grad()
        b.forward_kld(x).backward()
        before = ops.LAUNCHES
        opt_b.step()
        assert ops.LAUNCHES == before + 1
    tol = 2e-5 if dtype == torch.float32 else 1e-10
    for (n, pa), pb in zip(a.named_parameters(), b.parameters()):
        assert torch.allclose(pa, pb, rtol=tol, atol=tol), (n, float((pa - pb).abs().max()))


@pytest.mark.parametrize("C,H,dtype", [(12, 16, torch.float32), (48, 4, torch.float32), (24, 8, torch.float64)])
def test_glow_block_folded_actnorm_conv_equals_the_separate_layers(C, H, dtype, monkeypatch):
    """GlowBlock.inverse with ActNorm.inverse + Invertible1x1Conv.inverse folded into one node (ops.glow_actnorm_conv)
    against the same block running its three layers one by one: values, log-det and every gradient."""
    from vcnf_b200.flows.affine.glow import GlowBlock
    torch.manual_seed(C)
    blk = GlowBlock(C, 64, scale_map="sigmoid").to(dtype).cuda()
    with torch.no_grad():
        for p in blk.parameters():
            p.add_(0.05 * torch.randn_like(p))
    z = torch.randn(6, C, H, H, generator=_g(11)).to(dtype).cuda()
    with torch.no_grad():
        blk.inverse(z)                                      # data-dependent ActNorm initialisation (separate layers)
    ref = copy.deepcopy(blk)
    monkeypatch.setattr(ref, "_fused_tail", lambda z: None)
    gz = torch.randn(6, C, H, H, generator=_g(12)).to(dtype).cuda()
    gl = torch.randn(6, generator=_g(13)).to(dtype).cuda()
    outs = []
    for m in (blk, ref):
        zi = z.clone().requires_grad_(True)
        before = ops.LAUNCHES
        y, ld = m.inverse(zi)
        ((y * gz).sum() + (ld * gl).sum()).backward()
        outs.append((y.detach(), ld.detach(), zi.grad, ops.LAUNCHES - before))
    assert outs[0][3] < outs[1][3]                          # the folded path really ran (fewer package launches)
    tol = 2e-5 if dtype == torch.float32 else 1e-11
    rel = lambda a, b: float((a - b).abs().max() / (b.abs().max() + 1e-30))     # noqa: E731
    assert rel(outs[0][0], outs[1][0]) <= tol and rel(outs[0][1], outs[1][1]) <= tol and rel(outs[0][2], outs[1][2]) <= tol
    for (n, pa), pb in zip(blk.named_parameters(), ref.parameters()):
        assert pa.grad is not None and pb.grad is not None, n
        assert rel(pa.grad, pb.grad) <= 10 * tol, (n, rel(pa.grad, pb.grad))


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
def test_batched_lu_weights_equal_the_per_layer_launches(dtype, monkeypatch):
    """forward_kld / log_prob with the LU weights of all layers formed by one launch and their log-dets absorbed into one
    model-level sum (flows/mixing.py precompute_lu_weights) against the same model forming them layer by layer."""
    torch.manual_seed(5)
    flows = []
    for _ in range(4):
        flows += [nf.flows.CoupledRationalQuadraticSpline(7, 1, 16), nf.flows.LULinearPermute(7, identity_init=False)]
    a = nf.NormalizingFlow(nf.distributions.DiagGaussian(7, trainable=False), flows).to(dtype).cuda()
    b = copy.deepcopy(a)
    monkeypatch.setattr(b, "_lu_prepare", lambda absorb=False: ([], None))
    x = torch.randn(777, 7, generator=_g(21)).to(dtype).cuda()
    before = ops.LAUNCHES
    la = a.forward_kld(x.clone())
    la.backward()
    na = ops.LAUNCHES - before
    before = ops.LAUNCHES
    lb = b.forward_kld(x.clone())
    lb.backward()
    nb = ops.LAUNCHES - before
    assert na == nb - 2 * (4 - 1)                              # 4 + 4 single-CTA launches became 1 + 1
    tol = 1e-5 if dtype == torch.float32 else 1e-12
    assert abs(float(la) - float(lb)) <= tol * max(1.0, abs(float(lb)))
    for (n, pa), pb in zip(a.named_parameters(), b.parameters()):
        if pb.grad is None:
            assert pa.grad is None, n
            continue
        d = float((pa.grad - pb.grad).abs().max())
        assert d <= 20 * tol * max(1.0, float(pb.grad.abs().max())), (n, d)
    with torch.no_grad():
        assert float((a.log_prob(x) - b.log_prob(x)).abs().max()) <= 10 * tol * 50
    assert all("_pre" not in m.__dict__ for m in a.modules())
