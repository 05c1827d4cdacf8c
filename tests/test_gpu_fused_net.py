"""vcnf_resnet_spline_fwd (the whole ResidualNet conditioner + the coupling spline in ONE launch, used by the no-grad
directions) against the layer-by-layer kernels of the same layer (which the parity suites pin to the oracle) and against
the oracle itself: both directions, ragged / multi-tile batches, every tail mode, with and without the unconditional
spline on the identity columns, 0-3 residual blocks."""
import pytest
import torch

import oracle
import vcnf_b200 as nf
from vcnf_b200 import ops
from vcnf_b200.flows.neural_spline.coupling import PiecewiseRationalQuadraticCoupling
from vcnf_b200.nets.resnet import ResidualNet
from vcnf_b200.utils.masks import create_alternating_binary_mask
from cases import randomize
from parity import assert_close, SPLINE

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _layerwise(fn, *args):
    ops.FUSED_NET = False
    try:
        return fn(*args)
    finally:
        ops.FUSED_NET = True


def _both_paths(layer, x, inverse):
    fn = layer.inverse if inverse else layer.forward
    with torch.no_grad():
        y_ref, ld_ref = _layerwise(fn, x)       # layer-by-layer kernels
        before = ops.LAUNCHES
        y, ld = fn(x)
    assert ops.LAUNCHES - before == 2           # pack + the fused kernel
    return y, ld, y_ref, ld_ref


@pytest.mark.parametrize("B", [1, 127, 128, 300, 20011])
@pytest.mark.parametrize("inverse", [False, True])
def test_cfg2_layer_fused_matches_layerwise(B, inverse):
    torch.manual_seed(B)
    layer = nf.flows.CoupledRationalQuadraticSpline(29, 2, 128)
    randomize(layer, scale=0.05)
    layer = layer.to(DEV)
    x = 2.0 * torch.randn(B, 29, device=DEV)    # tail bound 3: a few per cent of the elements fall in the tails
    y, ld, y_ref, ld_ref = _both_paths(layer, x, inverse)
    assert_close(y, y_ref, **SPLINE, what="fused cfg2 layer y")
    assert_close(ld, ld_ref, rtol=SPLINE["rtol"], atol=29 * SPLINE["atol"], what="fused cfg2 layer logdet")


@pytest.mark.parametrize("tails,uncond,blocks,D", [("linear", False, 1, 6), ("circular", True, 2, 9), (None, False, 0, 5),
                                                   (["linear", "circular"] * 8, True, 3, 16), ("linear", True, 2, 32)])
def test_fused_variants(tails, uncond, blocks, D):
    torch.manual_seed(D)
    mask = create_alternating_binary_mask(D, even=False)
    layer = PiecewiseRationalQuadraticCoupling(
        mask, lambda i, o: ResidualNet(i, o, hidden_features=128, num_blocks=blocks), num_bins=8, tails=tails,
        tail_bound=2.5, apply_unconditional_transform=uncond)
    randomize(layer, scale=0.05)
    layer = layer.to(DEV)
    x = torch.rand(777, D, device=DEV) if tails is None else 1.5 * torch.randn(777, D, device=DEV)
    for inverse in (False, True):
        y, ld, y_ref, ld_ref = _both_paths(layer, x, inverse)
        assert_close(y, y_ref, **SPLINE, what="fused variant y")
        assert_close(ld, ld_ref, rtol=SPLINE["rtol"], atol=D * SPLINE["atol"], what="fused variant logdet")


def test_fused_outputs_params_and_activations():
    """The optional outputs: spline parameters and the pre-activation output of every linear layer."""
    torch.manual_seed(5)
    layer = nf.flows.CoupledRationalQuadraticSpline(29, 2, 128)
    randomize(layer, scale=0.05)
    layer = layer.to(DEV)
    c = layer.prqct
    x = torch.randn(1000, 29, device=DEV)
    with torch.no_grad():
        fused = c._fused_route(x, None)
        assert fused is not None
        y, ld, params, acts = ops.resnet_spline(x, fused, *c.transform_net.fused_params(), inverse=False,
                                                shared=c._shared_params(), want_params=True, want_acts=True)
        net = c.transform_net
        xin = x[:, c.identity_features.to(DEV)]
        h = torch.nn.functional.linear(xin.double(), net.initial_layer.weight.double(), net.initial_layer.bias.double())
        ref = [h]
        for blk in net.blocks:
            l0, l1 = blk.linear_layers
            t = torch.nn.functional.linear(torch.relu(h), l0.weight.double(), l0.bias.double())
            h = h + torch.nn.functional.linear(torch.relu(t), l1.weight.double(), l1.bias.double())
            ref += [t, h]
        theta = torch.nn.functional.linear(h, net.final_layer.weight.double(), net.final_layer.bias.double())
    for i, r in enumerate(ref):
        assert_close(acts[i], r.float(), rtol=1e-5, atol=1e-5 * float(r.abs().max()), what="fused acts")
    assert_close(params, theta.float(), rtol=1e-5, atol=1e-5 * float(theta.abs().max()), what="fused params")


def test_fused_layer_against_oracle():
    torch.manual_seed(11)
    ref = oracle.flows.CoupledRationalQuadraticSpline(29, 2, 128)
    randomize(ref, scale=0.05)
    layer = nf.flows.CoupledRationalQuadraticSpline(29, 2, 128)
    layer.load_state_dict(ref.state_dict())
    layer = layer.to(DEV)
    x = 1.5 * torch.randn(513, 29)
    ref = ref.double()
    with torch.no_grad():
        for inverse in (False, True):
            yr, lr = (ref.inverse if inverse else ref.forward)(x.double())
            y, ld = (layer.inverse if inverse else layer.forward)(x.to(DEV))
            assert_close(y.cpu(), yr.float(), **SPLINE, what="fused vs oracle y")
            assert_close(ld.cpu(), lr.float(), rtol=SPLINE["rtol"], atol=29 * SPLINE["atol"], what="fused vs oracle logdet")


@pytest.mark.parametrize("B", [300, 4099])
def test_training_route_values_and_gradients(B):
    """With gradients the density direction takes its VALUES from the fused launch and builds the layer-by-layer autograd
    nodes around them: outputs and every gradient against the layer-by-layer path."""
    torch.manual_seed(B)
    layer = nf.flows.CoupledRationalQuadraticSpline(29, 2, 128)
    randomize(layer, scale=0.05)
    layer = layer.to(DEV)
    x = 1.5 * torch.randn(B, 29, device=DEV)
    gz = torch.randn(B, 29, device=DEV)
    gl = torch.randn(B, device=DEV)

    def run():
        xi = x.clone().requires_grad_(True)
        layer.zero_grad(set_to_none=True)
        z, ld = layer.inverse(xi)               # density direction (prqct.forward)
        ((z * gz).sum() + (ld * gl).sum()).backward()
        return z.detach(), ld.detach(), xi.grad, {n: p.grad.clone() for n, p in layer.named_parameters()}

    before = ops.LAUNCHES
    z, ld, gx, gp = run()
    fused_launches = ops.LAUNCHES - before
    before = ops.LAUNCHES
    zr, lr, gxr, gpr = _layerwise(run)
    assert fused_launches < ops.LAUNCHES - before
    assert_close(z, zr, **SPLINE, what="fused train z")
    assert_close(ld, lr, rtol=SPLINE["rtol"], atol=29 * SPLINE["atol"], what="fused train logdet")
    assert_close(gx, gxr, rtol=1e-4, atol=1e-4 * float(gxr.abs().max()), what="fused train gx")
    for n, r in gpr.items():
        assert_close(gp[n], r, rtol=1e-4, atol=1e-4 * float(r.abs().max()), what="fused train grad " + n)


def test_c_abi_padded_rows_and_accumulate():
    """Straight through the C ABI: x and y as [B, 29] views of wider storage (row pitch 32: the element-wise tile copies),
    log-det accumulated onto an existing vector, and an empty batch."""
    import ctypes
    from vcnf_b200 import _lib
    from vcnf_b200._lib import ptr, stream
    torch.manual_seed(3)
    layer = nf.flows.CoupledRationalQuadraticSpline(29, 2, 128)
    randomize(layer, scale=0.05)
    layer = layer.to(DEV)
    c = layer.prqct
    B = 1000
    x = 1.5 * torch.randn(B, 29, device=DEV)
    with torch.no_grad():
        y_ref, ld_ref = c.forward(x)
        f = c._fused_route(x, None)
        w, b = c.transform_net.fused_params()
        packed = f.pack(w, b)
        shared = c._shared_params().contiguous()
        xp = torch.zeros(B, 32, device=DEV)
        xp[:, :29] = x
        yp = torch.full((B, 32), 7.0, device=DEV)
        ld = torch.full((B,), 2.0, device=DEV)
        err = torch.zeros(1, dtype=torch.int32, device=DEV)

        def call(rows):
            _lib.call("vcnf_resnet_spline_fwd", ptr(xp), 32, ptr(yp), 32, ptr(ld), 1, rows, 29, 8, f.id_cols, f.n_in,
                      f.tr_cols, f.Tc, f.Pc, f.hidden, f.num_blocks, ptr(packed), f.modes, f.bounds, f.wh_scale, 0,
                      f.outside_zero, ptr(err), ptr(shared), f.Ts, f.Ps, f.smodes, f.sbounds, None, 0, None, None, 0,
                      stream())
        call(0)                                    # nothing is touched
        assert float(yp.min()) == 7.0 and float(ld.min()) == 2.0
        call(B)
    assert torch.equal(yp[:, :29], y_ref) and float(yp[:, 29:].min()) == 7.0     # same kernel, same tile arithmetic
    assert_close(ld, ld_ref + 2.0, rtol=1e-6, atol=1e-5, what="fused accumulate logdet")
    assert int(err) == 0


def test_unsupported_shapes_keep_the_layerwise_kernels():
    """hidden != 128 is not covered: the layer runs layer by layer, with the same results API."""
    torch.manual_seed(4)
    layer = nf.flows.CoupledRationalQuadraticSpline(8, 1, 64).to(DEV)
    x = torch.randn(100, 8, device=DEV)
    with torch.no_grad():
        assert layer.prqct._fused_route(x, None) is None
        before = ops.LAUNCHES
        layer.forward(x)
        assert ops.LAUNCHES - before > 3


def test_fused_training_step_against_oracle_gradients():
    """forward_kld of a two-layer cfg-2-shaped model (hidden 128: the fused forward with saved outputs + the
    layer-by-layer backward kernels) against the oracle evaluated in fp64: loss and every parameter gradient."""
    from cases import build_cfg2_small
    from parity import assert_grads_close
    torch.manual_seed(21)
    ref = build_cfg2_small(oracle, layers=2, d=29, hidden=128)
    randomize(ref, scale=0.05)
    model = build_cfg2_small(nf, layers=2, d=29, hidden=128)
    model.load_state_dict(ref.state_dict())
    model = model.to(DEV)
    ref = ref.double()
    x = 1.2 * torch.randn(512, 29)
    before = ops.LAUNCHES
    loss = model.forward_kld(x.to(DEV))
    loss.backward()
    fused_launches = ops.LAUNCHES - before
    lr = -torch.mean(ref.log_prob(x.double()))          # = forward_kld without categorical columns (core.py:36-67)
    lr.backward()
    assert_close(loss.detach().cpu(), lr.detach().float(), rtol=1e-4, atol=2e-4, what="fused train loss vs oracle")
    got = {n: p.grad.detach().cpu() for n, p in model.named_parameters()}
    want = {n: (p.grad.float() if p.grad is not None else None) for n, p in ref.named_parameters()}
    assert_grads_close(got, want, rtol=1e-4, what="fused train vs oracle")
    model.zero_grad(set_to_none=True)
    before = ops.LAUNCHES
    _layerwise(lambda: model.forward_kld(x.to(DEV)).backward())
    assert fused_launches < ops.LAUNCHES - before          # the fused route really ran


def test_sample_with_batched_lu_inverses_matches_the_per_layer_path():
    """sample() without gradients inverts all LU layers in one launch and folds their log-dets into one term; with
    gradients enabled every layer forms its own inverse (and the coupling layers run layer by layer)."""
    from cases import build_cfg2_small
    torch.manual_seed(31)
    model = build_cfg2_small(nf, layers=3, d=29, hidden=128)
    randomize(model, scale=0.05)
    model = model.to(DEV)
    torch.manual_seed(5)
    with torch.no_grad():
        before = ops.LAUNCHES
        z, lq = model.sample(777)
        fast = ops.LAUNCHES - before
    torch.manual_seed(5)
    before = ops.LAUNCHES
    zr, lqr = _layerwise(model.sample, 777)     # grad mode: per-layer inverses, layer-by-layer coupling kernels
    assert fast < ops.LAUNCHES - before
    assert_close(z, zr.detach(), **SPLINE, what="sample z, batched LU inverses")
    assert_close(lq, lqr.detach(), rtol=SPLINE["rtol"], atol=3 * 29 * SPLINE["atol"], what="sample log_q, batched LU inverses")
    for f in model.flows:                       # nothing is left behind for a later call
        if hasattr(f, "linear"):
            assert "_pre_inv" not in f.linear.__dict__


def test_graphed_log_prob_matches_eager_and_sees_new_parameters():
    """parallel.GraphedLogProb: the replayed evaluation equals the eager one, follows the input it is given, and
    (weights are re-tiled inside the graph) follows parameter updates made between replays."""
    from vcnf_b200 import parallel
    from cases import build_cfg2_small
    torch.manual_seed(41)
    model = build_cfg2_small(nf, layers=2, d=29, hidden=128)
    randomize(model, scale=0.05)
    model = model.to(DEV)
    x1, x2 = torch.randn(2048, 29, device=DEV), 1.3 * torch.randn(2048, 29, device=DEV)
    lp = parallel.GraphedLogProb(model, x1)
    assert lp.graph is not None, lp.capture_error
    with torch.no_grad():
        for x in (x1, x2):
            assert torch.equal(lp(x), model.log_prob(x))
        for p in model.parameters():
            p.add_(0.01 * torch.randn_like(p))
        assert torch.equal(lp(x2), model.log_prob(x2))
