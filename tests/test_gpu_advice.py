"""Regression tests for the round-1 advisor findings: non-finite loss must not reach the parameters or the moments,
the graphed step's warm-up must not train, per-row linear maps wider than 64, MADE's in-step masked weight under
programmatic dependent launch, data-parallel start-up (same weights on every rank, equal shards) and a 2-rank NCCL step
against one rank on the concatenated batch."""
import copy
import os

import pytest
import torch

import oracle
import vcnf_b200 as nf
from vcnf_b200 import optim, parallel
from cases import build_cfg2_small, randomize
from harness import eval_layer
from parity import assert_close, assert_grads_close

pytestmark = pytest.mark.gpu


def _small_model(seed=0):
    torch.manual_seed(seed)
    m = build_cfg2_small(nf, layers=2, d=6, hidden=16)
    randomize(m, scale=0.05)
    return m.cuda()


def test_non_finite_loss_skips_the_whole_update():
    model = _small_model()
    opt = optim.FlatAdam(model.parameters(), lr=1e-2, weight_decay=1e-3)
    x = torch.randn(256, 6, device="cuda")
    for _ in range(2):                                   # two ordinary steps so the moments are non-zero
        opt.zero_grad()
        loss = model.forward_kld(x)
        loss.backward()
        opt.guard = loss.detach()
        opt.step()
    p0, m0, v0, s0 = opt.flat_p.clone(), opt.exp_avg.clone(), opt.exp_avg_sq.clone(), opt.steps.clone()
    opt.zero_grad()
    loss = model.forward_kld(x)
    loss.backward()
    for bad in (float("nan"), float("inf"), -float("inf")):
        opt.guard = torch.full((), bad, device="cuda")
        opt.step()
        torch.cuda.synchronize()
        assert torch.equal(opt.flat_p, p0) and torch.equal(opt.exp_avg, m0) and torch.equal(opt.exp_avg_sq, v0)
        assert torch.equal(opt.steps, s0)
    opt.guard = loss.detach()                            # a finite loss steps again
    opt.step()
    assert not torch.equal(opt.flat_p, p0) and float(opt.steps) == float(s0) + 1


def test_graphed_step_warmup_does_not_train_and_lr_edits_reach_the_replay():
    model = _small_model()
    opt = optim.FlatAdam(model.parameters(), lr=1e-2)
    x = torch.randn(512, 6, device="cuda")
    before = opt.flat_p.clone()
    step = parallel.GraphedTrainStep(model, opt, opt.bucket, x, world=1)
    assert step.graph is not None, step.capture_error
    assert torch.equal(opt.flat_p, before) and float(opt.steps) == 0.0 and float(opt.exp_avg.abs().max()) == 0.0
    step(x)
    after_one = opt.flat_p.clone()
    assert float(opt.steps) == 1.0 and not torch.equal(after_one, before)
    opt.param_groups[0]["lr"] = 0.0                      # edited between replays: the kernel reads the device scalar
    step(x)
    assert float(opt.steps) == 2.0 and torch.equal(opt.flat_p, after_one)
    # a NaN batch: the replayed step leaves everything as it was
    m2 = opt.exp_avg.clone()
    step(torch.full_like(x, float("nan")))
    torch.cuda.synchronize()
    assert float(opt.steps) == 2.0 and torch.equal(opt.flat_p, after_one) and torch.equal(opt.exp_avg, m2)
    # library optimiser: state restored after the lr = 0 warm-up
    model2 = _small_model(1)
    opt2 = torch.optim.Adam(model2.parameters(), lr=1e-2, capturable=True)
    w0 = [p.detach().clone() for p in model2.parameters()]
    bucket2 = parallel.GradBucket(model2.parameters())
    parallel.GraphedTrainStep(model2, opt2, bucket2, x, world=1)
    assert all(torch.equal(a, b) for a, b in zip(w0, model2.parameters()))
    assert all(float(s.get("step", 0)) == 0 for s in opt2.state.values())
    assert opt2.param_groups[0]["lr"] == 1e-2


@pytest.mark.parametrize("layer", ["lu", "inv1x1", "inv1x1_lu"])
def test_linear_maps_wider_than_64(layer):
    D = 96
    if layer == "lu":
        build = lambda m: m.flows.LULinearPermute(D, identity_init=False)
        x = torch.randn(300, D, generator=torch.Generator().manual_seed(1))
    else:
        build = lambda m: m.flows.Invertible1x1Conv(D, use_lu=layer == "inv1x1_lu")
        x = torch.randn(3, D, 4, 4, generator=torch.Generator().manual_seed(2))
    torch.manual_seed(0)
    ref = build(oracle)
    randomize(ref, scale=0.02)
    torch.manual_seed(0)
    prod = build(nf)
    prod.load_state_dict(ref.state_dict())
    prod = prod.cuda()
    for direction in ("inverse", "forward"):
        r = eval_layer(copy.deepcopy(ref), x, direction)
        g = eval_layer(copy.deepcopy(prod), x.cuda(), direction)
        scale = float(r["z"].abs().max())
        assert_close(g["z"], r["z"], rtol=1e-5, atol=1e-5 * scale, what="%s96.z.%s" % (layer, direction))
        assert_close(g["log_det"], r["log_det"], rtol=1e-5, atol=1e-4, what="%s96.ld.%s" % (layer, direction))
        assert_grads_close({"x": g["gx"], **g["grads"]}, {"x": r["gx"], **r["grads"]}, rtol=2e-4,
                           what="%s96.%s" % (layer, direction))


def test_made_masked_weight_is_ordered_under_dependent_launch():
    """MADE multiplies W by its mask in an ATen kernel right before the GEMM (nets/made.py); the GEMM must not
    fetch W ahead of its dependency wait.  Repeated evaluations against an fp64 product."""
    torch.manual_seed(0)
    made = nf.nets.MADE(64, 256, num_blocks=2, output_multiplier=2).cuda()
    x = torch.randn(4096, 64, device="cuda")
    ref = copy.deepcopy(made).double()
    want = ref(x.double())
    for _ in range(20):
        got = made(x)
        assert float((got.double() - want).abs().max()) < 1e-4 * float(want.abs().max())


def _dp_worker(rank, world, port, x_all, state, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      LOCAL_RANK=str(rank))
    parallel.init_from_env("nccl")
    torch.cuda.set_device(rank)
    torch.manual_seed(100 + rank)                        # ranks seeded differently on purpose: sync_model fixes it
    model = build_cfg2_small(nf, layers=2, d=6, hidden=16).cuda()
    if rank == 0:
        model.load_state_dict(state)
    opt = optim.FlatAdam(model.parameters(), lr=1e-2)
    lo, hi = parallel.shard_bounds(len(x_all), rank, world)
    xs = x_all[lo:hi].cuda()
    step = parallel.GraphedTrainStep(model, opt, opt.bucket, xs)
    for _ in range(3):
        step(xs)
    torch.cuda.synchronize()
    if rank == 0:
        out.put({k: v.cpu().numpy() for k, v in model.state_dict().items()})
    torch.distributed.barrier()
    torch.cuda.synchronize()
    out.close()
    out.join_thread()
    os._exit(0)        # tearing the communicator down under live CUDA graphs that captured its collectives can block


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_two_rank_nccl_step_equals_one_rank_on_the_concatenated_batch():
    import torch.multiprocessing as mp
    torch.manual_seed(0)
    ref_model = build_cfg2_small(nf, layers=2, d=6, hidden=16)
    randomize(ref_model, scale=0.05)
    state = copy.deepcopy(ref_model.state_dict())
    x_all = 1.5 * torch.randn(1024, 6, generator=torch.Generator().manual_seed(3))
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    procs = [ctx.Process(target=_dp_worker, args=(r, 2, 29611, x_all, state, out)) for r in range(2)]
    for p in procs:
        p.start()
    got = {k: torch.from_numpy(v) for k, v in out.get(timeout=120).items()}
    for p in procs:
        p.join(60)
        if p.exitcode is None:
            p.kill()
        assert p.exitcode == 0
    single = ref_model.cuda()
    opt = optim.FlatAdam(single.parameters(), lr=1e-2)
    step = parallel.GraphedTrainStep(single, opt, opt.bucket, x_all.cuda(), world=1)
    for _ in range(3):
        step(x_all.cuda())
    for k, v in single.state_dict().items():
        if v.dtype.is_floating_point:
            assert_close(got[k], v, rtol=1e-4, atol=1e-5, what="dp2." + k)


def test_loaded_tail_bound_buffer_rebuilds_the_kernel_spec():
    """A tensor `tail_bound` is a buffer; the spline layers cache it inside their kernel spec on first use, so a
    load_state_dict that changes it must invalidate the cache."""
    import vcnf_b200 as nf
    from vcnf_b200.flows.neural_spline.autoregressive import MaskedPiecewiseRationalQuadraticAutoregressive as AR
    torch.manual_seed(4)
    x = 2.5 * torch.randn(301, 6, device="cuda")

    def build(bound):
        torch.manual_seed(9)
        return AR(6, 32, num_bins=8, tails="linear", tail_bound=torch.tensor(bound), init_identity=False).cuda()
    a, b = build(1.0), build(4.0)
    with torch.no_grad():
        ya0, _ = a.forward(x)                                   # builds a's spec with bound 1
        a.load_state_dict(b.state_dict())
        ya, lda = a.forward(x)
        yb, ldb = b.forward(x)
        assert not torch.equal(ya0, yb)
        assert torch.equal(ya, yb) and torch.equal(lda, ldb)
        za, _ = a.inverse(yb)
        zb, _ = b.inverse(yb)
        assert torch.equal(za, zb)
