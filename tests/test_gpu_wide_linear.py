"""GPU parity of the host-side compositions on top of the tcgen05 dense-layer kernels: contraction
widths > 128 (MADE, hidden 256: nets/made.py:79-80) run as 128-wide slices that accumulate in place, and
the tiny MLP conditioners of cfg 1 / cfg 4 (nets/mlp.py:7-58) run layer by layer with the ReLU fused into
the next layer's operand load.  Checked against an fp64 torch evaluation of the same fp32 operands;
bar = an fp32 FMA chain, |err| <= 2e-6 * sum |terms| (values) and 1e-5 relative to the largest gradient."""
import pytest
import torch
from torch.nn import functional as F

import oracle
import vcnf_b200 as nf
from vcnf_b200 import ops

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.mark.parametrize("M,N,K,relu,res", [(2000, 256, 256, True, True), (1111, 1472, 256, False, False),
                                            (512, 64, 200, True, False), (300, 2, 32, True, False),
                                            (300, 32, 1, False, False)])
def test_sliced_linear_forward_and_gradients(M, N, K, relu, res):
    g = torch.Generator(device=DEV).manual_seed(M + N + K)
    x = torch.randn(M, K, device=DEV, generator=g, requires_grad=True)
    w = (0.2 * torch.randn(N, K, device=DEV, generator=g)).requires_grad_(True)
    b = torch.randn(N, device=DEV, generator=g, requires_grad=True)
    r = torch.randn(M, N, device=DEV, generator=g, requires_grad=True) if res else None
    gy = torch.randn(M, N, device=DEV, generator=g)
    assert ops.linear_supported(x, w)
    y = ops.linear(x, w, b, residual=r, relu_in=relu)
    grads = torch.autograd.grad(y, [x, w, b] + ([r] if res else []), gy)

    x64, w64, b64 = (t.detach().double().requires_grad_(True) for t in (x, w, b))
    r64 = r.detach().double().requires_grad_(True) if res else None
    a64 = F.relu(x64) if relu else x64
    y64 = F.linear(a64, w64, b64) + (r64 if res else 0)
    grads64 = torch.autograd.grad(y64, [x64, w64, b64] + ([r64] if res else []), gy.double())
    mag = a64.detach().abs() @ w64.detach().abs().t() + b64.detach().abs() + (r64.detach().abs() if res else 0)
    assert float(((y.double() - y64).abs() / mag).max()) <= 2e-6
    for a, e in zip(grads, grads64):
        assert float((a.double() - e).abs().max()) <= 1e-5 * float(e.abs().max()) + 1e-30


@pytest.mark.parametrize("layers", [[1, 32, 32, 2], [29, 58, 29], [15, 30, 15]])
def test_mlp_conditioner_matches_oracle(layers):
    torch.manual_seed(3)
    ref = oracle.nets.MLP(layers)
    net = nf.nets.MLP(layers)
    net.load_state_dict(ref.state_dict())
    net = net.to(DEV)
    x = torch.randn(777, layers[0])
    launches = ops.LAUNCHES
    y = net(x.to(DEV).requires_grad_(True))
    # widths <= 32: the whole net is one register-resident kernel; wider: one tcgen05 launch per dense layer
    assert ops.LAUNCHES - launches == (1 if max(layers) <= 32 else len(layers) - 1)
    y.sum().backward()
    yr = ref(x)
    yr.sum().backward()
    assert torch.allclose(y.cpu(), yr, rtol=1e-5, atol=1e-6)
    for (n, p), (_, q) in zip(net.named_parameters(), ref.named_parameters()):
        assert torch.allclose(p.grad.cpu(), q.grad, rtol=1e-4, atol=1e-4 * float(q.grad.abs().max())), n


def test_made_runs_on_the_tensor_core_kernels():
    torch.manual_seed(4)
    ref = oracle.nets.MADE(features=64, hidden_features=256, num_blocks=2, output_multiplier=23)
    net = nf.nets.MADE(features=64, hidden_features=256, num_blocks=2, output_multiplier=23)
    for p in ref.parameters():
        p.data.normal_(0, 0.05)
    net.load_state_dict(ref.state_dict())
    net = net.to(DEV)
    x = torch.randn(1000, 64)
    launches = ops.LAUNCHES
    y = net(x.to(DEV))
    assert ops.LAUNCHES - launches >= 6 + 5          # 6 masked layers; the 5 that read hidden 256 run two slices
    (y * y).sum().backward()
    yr = ref(x)
    (yr * yr).sum().backward()
    assert torch.allclose(y.cpu(), yr, rtol=1e-4, atol=1e-5)
    for (n, p), (_, q) in zip(net.named_parameters(), ref.named_parameters()):
        assert torch.allclose(p.grad.cpu(), q.grad, rtol=1e-3, atol=1e-4 * float(q.grad.abs().max())), n


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64])
@pytest.mark.parametrize("layers,leaky,B", [([1, 32, 32, 2], 0.0, 1000), ([15, 30, 15], 0.1, 333), ([1, 4, 1], 0.0, 65),
                                            ([7, 32, 32, 32, 5], 0.2, 4099), ([3, 16], 0.0, 17),
                                            # above kMlpQuadMaxRows (8192): one thread per sample instead of four lanes
                                            ([1, 32, 32, 2], 0.0, 20011), ([7, 32, 32, 32, 5], 0.2, 16385),
                                            ([15, 30, 15], 0.1, 8193), ([15, 30, 15], 0.1, 8192)])
def test_register_mlp_matches_torch(layers, leaky, B, dtype):
    """vcnf_mlp_fwd / vcnf_mlp_bwd (whole MLP in registers) against the oracle net in the same dtype: values, input
    gradient and every parameter gradient; ragged batches, 1-4 linear layers, leaky slopes."""
    torch.manual_seed(B)
    ref = oracle.nets.MLP(layers, leaky=leaky).to(dtype)
    net = nf.nets.MLP(layers, leaky=leaky).to(dtype)
    net.load_state_dict(ref.state_dict())
    net = net.to(DEV)
    x = torch.randn(B, layers[0], dtype=dtype)
    gy = torch.randn(B, layers[-1], dtype=dtype)
    xr = x.clone().requires_grad_(True)
    xg = x.to(DEV).requires_grad_(True)
    launches = ops.LAUNCHES
    y = net(xg)
    ws = [m.weight for m in net.net if isinstance(m, torch.nn.Linear)]
    # (four 32-wide fp64 layers exceed the backward's shared-memory accumulators: that one case keeps nn.Linear)
    assert ops.LAUNCHES - launches == (1 if ops.mlp_small_supported(xg, ws, leaky) else 0)
    assert ops.mlp_small_supported(xg, ws, leaky) or (dtype == torch.float64 and len(layers) == 5)
    y.backward(gy.to(DEV))
    yr = ref(xr)
    yr.backward(gy)
    rt, at = (1e-5, 1e-6) if dtype == torch.float32 else (1e-11, 1e-12)
    assert torch.allclose(y.detach().cpu(), yr.detach(), rtol=rt, atol=at)
    assert torch.allclose(xg.grad.cpu(), xr.grad, rtol=10 * rt, atol=10 * at * float(xr.grad.abs().max()))
    for (n, p), (_, q) in zip(net.named_parameters(), ref.named_parameters()):
        assert torch.allclose(p.grad.cpu(), q.grad, rtol=10 * rt, atol=20 * rt * float(q.grad.abs().max())), n


def test_register_mlp_empty_batch():
    net = nf.nets.MLP([1, 32, 32, 2]).to(DEV)
    x = torch.zeros(0, 1, device=DEV, requires_grad=True)
    y = net(x)
    y.sum().backward()
    assert y.shape == (0, 2) and all(float(p.grad.abs().max()) == 0.0 for p in net.parameters())
