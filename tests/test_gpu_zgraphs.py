"""cfg 1 (README 2-D Real NVP: 16 x [AffineCouplingBlock(MLP[1,32,32,2]), Permute swap], TwoMoons target,
reverse_kld -- README.md:60-84 of the reference) as ONE replayed CUDA graph, and what happens when capture
cannot be done.  Runs last (file name): the second test forces a failed capture on purpose."""
import pytest
import torch

import vcnf_b200 as nf
from vcnf_b200 import ops, parallel

pytestmark = pytest.mark.gpu


def _cfg1(blocks=16):
    flows = []
    for _ in range(blocks):
        flows += [nf.flows.AffineCouplingBlock(nf.nets.MLP([1, 32, 32, 2], init_zeros=False)),
                  nf.flows.Permute(2, mode="swap")]
    return nf.NormalizingFlow(nf.distributions.DiagGaussian(2), flows, nf.distributions.TwoMoons()).cuda()


def test_cfg1_reverse_kld_step_replays_from_a_graph():
    """The whole step (base draw, 16 coupling blocks through the register MLP kernels, target log-density,
    backward, Adam) captured once; each replay draws fresh base samples and moves the parameters, and the
    loss goes down over a few hundred replays as it does eagerly."""
    torch.manual_seed(0)
    model = _cfg1()
    B = 1024
    opt = torch.optim.Adam(model.parameters(), lr=1e-3, capturable=True)
    bucket = parallel.GradBucket(model.parameters())
    dummy = torch.zeros(1, device="cuda")
    step = parallel.GraphedTrainStep(model, opt, bucket, dummy, world=1,
                                     loss_fn=lambda m, x: m.reverse_kld(num_samples=B))
    assert step.graph is not None, step.capture_error
    before = [p.detach().clone() for p in model.parameters()]
    launches = ops.LAUNCHES
    losses = [float(step(dummy)) for _ in range(300)]
    assert ops.LAUNCHES == launches            # replays: no host-side launches of the library
    assert all(l == l and abs(l) < 1e6 for l in losses)
    assert len(set(losses[:10])) == 10         # fresh draws every replay
    assert any(not torch.equal(a, p.detach()) for a, p in zip(before, model.parameters()))
    first, last = sum(losses[:20]) / 20, sum(losses[-20:]) / 20
    assert last < first - 0.1, (first, last)
    # a replay and an eager evaluation compute the same loss for the same generator state and parameters
    state = torch.cuda.get_rng_state()
    with torch.no_grad():
        eager = float(model.reverse_kld(num_samples=B))
    torch.cuda.set_rng_state(state)
    replayed = float(step(dummy))
    assert abs(eager - replayed) <= 1e-4 * max(1.0, abs(eager)), (eager, replayed)


def test_failed_capture_falls_back_to_eager_steps():
    """A loss kept from an eager backward on the default stream pins the parameters' AccumulateGrad nodes to the
    legacy stream; capture then fails.  The step object must say so, leave the CUDA generator usable and run
    the same step eagerly."""
    torch.manual_seed(1)
    model = _cfg1(blocks=2)
    B = 256
    opt0 = torch.optim.Adam(model.parameters(), lr=1e-3)
    kept = model.reverse_kld(num_samples=B)
    kept.backward()
    opt0.step()
    torch.cuda.synchronize()
    opt = torch.optim.Adam(model.parameters(), lr=1e-3, capturable=True)
    bucket = parallel.GradBucket(model.parameters())
    dummy = torch.zeros(1, device="cuda")
    step = parallel.GraphedTrainStep(model, opt, bucket, dummy, world=1,
                                     loss_fn=lambda m, x: m.reverse_kld(num_samples=B))
    if step.graph is None:
        assert "hint" in step.capture_error
    a, b = float(step(dummy)), float(step(dummy))          # eager (or graphed) steps still work, RNG included
    assert a == a and b == b and a != b
    z, _ = model.sample(16)
    assert torch.isfinite(z).all()
    del kept
