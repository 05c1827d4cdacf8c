"""GPU parity of the conditioner dense layers THROUGH THE C ABI (vcnf_linear_tf32x3,
vcnf_linear_wgrad_tf32x3) against an fp64 torch product of the same fp32 operands.

The kernels split every fp32 operand into TF32 hi + lo and accumulate three products on the tensor
cores; the stated bar is an fp32 FMA chain: |err| <= 2e-6 * sum_k |a_k w_k| (cuBLAS fp32 measures
1.5-3e-7 of that sum on the same data, these kernels 2-4e-7).  Shapes cover every staging path: rows
with a 16-byte-multiple pitch (tensor-map TMA loads / stores), odd pitches (cp.async / scalar
fallback), ragged M, column slices (N > 128), the transposed weight walk of the input gradient,
ReLU on the input, bias, residual, ReLU-backward mask, in-place accumulation through `residual`."""
import pytest
import torch

from vcnf_b200 import _lib
from vcnf_b200._lib import ptr, stream

pytestmark = pytest.mark.gpu
DEV = "cuda"
BAR = 2e-6


def _padded(rows, cols, pitch):
    return torch.zeros(rows, pitch, device=DEV)[:, :cols]


def _fill(t, gen, scale=1.0):
    t.copy_(scale * torch.randn(t.shape, generator=gen, device=DEV))
    return t


CASES = [
    # M, N, K, lda, ldc, transposed W, relu, bias, residual, mask
    (128, 128, 32, 32, 128, False, False, False, False, False),
    (300, 128, 128, 128, 128, False, True, True, True, False),       # ragged M, TMA everywhere
    (1000, 128, 15, 15, 128, False, False, True, False, False),      # odd input pitch: fallback loads
    (1000, 128, 15, 16, 128, False, False, True, False, False),      # padded pitch: TMA loads, K tail
    (4096, 322, 128, 128, 322, False, False, True, False, False),    # 3 column slices, odd output pitch
    (4096, 322, 128, 128, 324, False, False, True, False, False),    # padded output pitch: TMA stores
    (777, 128, 128, 128, 128, True, False, False, True, True),       # input gradient: W^T, mask, residual
    (777, 15, 128, 128, 16, True, False, False, False, False),
    (65536, 128, 128, 128, 128, False, True, True, True, False),     # bench shape
    (5000, 64, 64, 64, 64, False, True, True, True, False),
]


@pytest.mark.parametrize("M,N,K,lda,ldc,wT,relu,bias,res,mask", CASES)
def test_linear_matches_fp64_product(M, N, K, lda, ldc, wT, relu, bias, res, mask):
    g = torch.Generator(device=DEV).manual_seed(M * 131 + N * 7 + K)
    A = _fill(_padded(M, K, lda), g)
    W = _fill(torch.empty((K, N) if wT else (N, K), device=DEV), g, 0.3)
    b = _fill(torch.empty(N, device=DEV), g) if bias else None
    R = _fill(_padded(M, N, ldc), g) if res else None
    Mk = _fill(_padded(M, N, ldc), g) if mask else None
    C = _padded(M, N, ldc)
    sw_n, sw_k = (1, N) if wT else (K, 1)
    _lib.call("vcnf_linear_tf32x3", ptr(A), lda, ptr(W), sw_n, sw_k, ptr(b), ptr(R), ldc if res else 0,
              ptr(Mk), ldc if mask else 0, ptr(C), ldc, M, N, K, int(relu), stream())
    a64 = (A.clamp_min(0) if relu else A).double()
    w64 = (W.t() if wT else W).double()
    ref = a64 @ w64.t()
    mag = a64.abs() @ w64.abs().t()
    if bias:
        ref, mag = ref + b.double(), mag + b.double().abs()
    if mask:
        ref = torch.where(Mk > 0, ref, torch.zeros_like(ref))
    if res:
        ref, mag = ref + R.double(), mag + R.double().abs()
    err = (C.double() - ref).abs() / mag.clamp_min(1e-30)
    assert torch.isfinite(C).all()
    assert float(err.max()) <= BAR, "max |err| / sum|terms| = %.3g" % float(err.max())


def test_linear_accumulates_in_place_through_residual():
    """The sliced input gradient of wide layers adds into its own output (residual == C)."""
    g = torch.Generator(device=DEV).manual_seed(5)
    M, N, K = 3000, 128, 66
    A = _fill(torch.empty(M, K, device=DEV), g)
    W = _fill(torch.empty(N, K, device=DEV), g, 0.3)
    C = _fill(torch.empty(M, N, device=DEV), g)
    before = C.double().clone()
    _lib.call("vcnf_linear_tf32x3", ptr(A), K, ptr(W), K, 1, None, ptr(C), N, None, 0, ptr(C), N, M, N, K, 0, stream())
    ref = before + A.double() @ W.double().t()
    mag = before.abs() + A.double().abs() @ W.double().abs().t()
    assert float(((C.double() - ref).abs() / mag).max()) <= BAR


WCASES = [
    # B, N, K, ldg, ldx, relu
    (32, 128, 128, 128, 128, False),
    (1000, 128, 128, 128, 128, True),
    (777, 128, 15, 128, 15, False),       # odd pitch: register-transposing fallback
    (777, 128, 15, 128, 16, False),       # padded pitch: TMA, ones column inside the data group
    (5000, 322, 128, 322, 128, False),    # odd pitch of gY
    (5000, 322, 128, 324, 128, False),    # padded: TMA, 3 output tiles
    (65536, 128, 128, 128, 128, True),    # bench shape
    (4097, 29, 29, 29, 29, False),        # LU linear weight gradient
]


@pytest.mark.parametrize("B,N,K,ldg,ldx,relu", WCASES)
def test_weight_gradient_matches_fp64_product(B, N, K, ldg, ldx, relu):
    g = torch.Generator(device=DEV).manual_seed(B + 3 * N + 5 * K)
    G = _fill(_padded(B, N, ldg), g)
    X = _fill(_padded(B, K, ldx), g)
    gW = torch.empty(N, K, device=DEV)
    gb = torch.empty(N, device=DEV)
    ws = torch.empty(_lib.query("vcnf_linear_wgrad_workspace", B, N, K), dtype=torch.uint8, device=DEV)
    _lib.call("vcnf_linear_wgrad_tf32x3", ptr(G), ldg, ptr(X), ldx, ptr(gW), ptr(gb), ptr(ws), B, N, K, int(relu), stream())
    x64 = (X.clamp_min(0) if relu else X).double()
    ref_w = G.double().t() @ x64
    mag_w = G.double().abs().t() @ x64.abs()
    ref_b = G.double().sum(0)
    mag_b = G.double().abs().sum(0)
    assert float(((gW.double() - ref_w).abs() / mag_w.clamp_min(1e-30)).max()) <= BAR
    assert float(((gb.double() - ref_b).abs() / mag_b.clamp_min(1e-30)).max()) <= BAR
    # fixed-order split-K reduction: bit-reproducible
    gW2 = torch.empty_like(gW)
    _lib.call("vcnf_linear_wgrad_tf32x3", ptr(G), ldg, ptr(X), ldx, ptr(gW2), ptr(gb), ptr(ws), B, N, K, int(relu), stream())
    assert torch.equal(gW, gW2)


def test_padded_pitch_layer_equals_contiguous_layer():
    """The padded-pitch plumbing of the coupling layer (ops.gather_cols(pad) -> ResidualNet(padded_out)
    -> spline kernel with ld_cond) gives bit-identical results to the contiguous route."""
    import vcnf_b200 as nf
    torch.manual_seed(3)
    layer = nf.flows.CoupledRationalQuadraticSpline(29, 2, 128).to(DEV)
    x = torch.randn(5000, 29, device=DEV)
    net = layer.prqct.transform_net
    assert layer.prqct._pad_rows and net.padded_out
    y1, ld1 = layer(x)
    layer.prqct._pad_rows, net.padded_out = False, False
    y2, ld2 = layer(x)
    assert torch.equal(y1, y2) and torch.equal(ld1, ld2)


@pytest.mark.parametrize("D,dtype", [(29, torch.float32), (29, torch.float64), (2, torch.float32), (64, torch.float64), (1, torch.float32)])
def test_lu_compose_matches_index_put_route(D, dtype):
    """Fused W = L U / log|det| kernel (vcnf_lu_compose_*) against the reference's construction
    (mixing.py:352-396: index_put, softplus, matmul, log-sum) in both directions of autograd."""
    from vcnf_b200 import ops
    from vcnf_b200.flows.mixing import _LULinear
    torch.manual_seed(D)
    lin = _LULinear(D, identity_init=False).to(DEV, dtype)
    with torch.no_grad():
        lin.unconstrained_upper_diag.add_(torch.randn(D, device=DEV, dtype=dtype))
    gW = torch.randn(D, D, device=DEV, dtype=dtype)
    gl = torch.randn((), device=DEV, dtype=dtype)
    params = [lin.lower_entries, lin.upper_entries, lin.unconstrained_upper_diag]
    W0, l0 = lin.weight(), lin.logabsdet()
    ref = torch.autograd.grad([W0, l0], params, [gW, gl], allow_unused=True)
    W1, l1 = ops.lu_compose(*params, lin.eps)
    got = torch.autograd.grad([W1, l1], params, [gW, gl], allow_unused=True)
    tol = dict(rtol=1e-5, atol=1e-6) if dtype == torch.float32 else dict(rtol=1e-12, atol=1e-13)
    torch.testing.assert_close(W1, W0, **tol)
    torch.testing.assert_close(l1, l0, **tol)
    for a, b in zip(got, ref):
        if b is not None and b.numel():
            torch.testing.assert_close(a, b, **tol)


def test_dependent_launch_chain_is_race_free():
    """The GEMM kernels are launched with programmatic stream serialization and signal their dependents early
    (gemm_tf32x3.cu, `pdl_trigger` / `pdl_wait`): a kernel's prologue may run while its predecessor drains.  A chain
    of dependent layers launched back to back must give bit-identical values and gradients to the same chain with
    a device synchronisation between all launches (which rules every overlap out), run after run."""
    from vcnf_b200 import ops
    torch.manual_seed(11)
    B, H = 40000, 128                               # 625 tiles: several per CTA, ragged last wave
    x0 = torch.randn(B, H, device="cuda")
    Ws = [torch.randn(H, H, device="cuda") * 0.08 for _ in range(6)]
    bs = [torch.randn(H, device="cuda") * 0.1 for _ in range(6)]

    def run(sync):
        x = x0.clone().requires_grad_(True)
        ws = [w.clone().requires_grad_(True) for w in Ws]
        bb = [b.clone().requires_grad_(True) for b in bs]

        def barrier(t):
            if sync:
                torch.cuda.synchronize()
                if t.requires_grad:
                    t.register_hook(lambda g: (torch.cuda.synchronize(), g)[1])
            return t
        h = barrier(ops.linear(x, ws[0], bb[0]))
        for k in (1, 3):                            # two residual blocks out of single launches
            t = barrier(ops.linear(h, ws[k], bb[k], relu_in=True))
            h = barrier(ops.linear(t, ws[k + 1], bb[k + 1], residual=h, relu_in=True))
        y = barrier(ops.linear(h, ws[5], bb[5], relu_in=True))
        gy = torch.ones_like(y) * 0.01
        grads = torch.autograd.grad(y, [x] + ws + bb, gy)
        torch.cuda.synchronize()
        return [y.detach()] + [g.detach() for g in grads]

    ref = run(sync=True)
    for _ in range(12):
        got = run(sync=False)
        for i, (a, b) in enumerate(zip(ref, got)):
            assert torch.equal(a, b), (i, float((a - b).abs().max()))
