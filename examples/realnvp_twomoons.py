"""The reference README's example (README.md:60-110: a 2-D Real NVP fitted to TwoMoons by reverse KL) with the import
swapped -- `import vcnf_b200 as nf` instead of `import normflow as nf` -- and, optionally, the whole step replayed from
a CUDA graph (`--graph`).  Prints the loss every 100 iterations; it falls from ~18 towards ~1.

    python examples/realnvp_twomoons.py [--iters 1000] [--graph]
"""
import argparse
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch                      # noqa: E402
import vcnf_b200 as nf            # noqa: E402   (the reference: import normflow as nf)

ap = argparse.ArgumentParser()
ap.add_argument("--iters", type=int, default=1000)
ap.add_argument("--graph", action="store_true")
args = ap.parse_args()

torch.manual_seed(0)
base = nf.distributions.base.DiagGaussian(2)
flows = []
for i in range(16):
    param_map = nf.nets.MLP([1, 32, 32, 2], init_zeros=True)
    flows.append(nf.flows.AffineCouplingBlock(param_map))
    flows.append(nf.flows.Permute(2, mode="swap"))
target = nf.distributions.target.TwoMoons()
model = nf.NormalizingFlow(base, flows, target).to("cuda")

if args.graph:
    from vcnf_b200 import parallel
    optimizer = nf.optim.FlatAdam(model.parameters(), lr=1e-3, weight_decay=1e-5)
    step = parallel.GraphedTrainStep(model, optimizer, optimizer.bucket, torch.zeros(1, device="cuda"),
                                     loss_fn=lambda m, x: m.reverse_kld(num_samples=1024))
    x0 = torch.zeros(1, device="cuda")
else:
    optimizer = torch.optim.Adam(model.parameters(), lr=1e-3, weight_decay=1e-5)

t0 = time.perf_counter()
for it in range(args.iters):
    if args.graph:
        loss = step(x0)
    else:
        optimizer.zero_grad()
        loss = model.reverse_kld(num_samples=1024)
        if not (torch.isnan(loss) | torch.isinf(loss)):
            loss.backward()
            optimizer.step()
    if it % 100 == 0 or it == args.iters - 1:
        print("iteration %4d  reverse KLD %.4f" % (it, float(loss.detach())))
torch.cuda.synchronize()
print("%.1f iterations/s" % (args.iters / (time.perf_counter() - t0)))
z, log_q = model.sample(4096)
print("sample: mean |x| = %.3f, finite = %s" % (float(z.abs().mean()), bool(torch.isfinite(z).all())))
