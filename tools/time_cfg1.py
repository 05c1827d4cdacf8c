"""GPU-box timing of cfg 1 (README 2-D Real NVP: 16 x [AffineCouplingBlock(MLP[1,32,32,2]), Permute swap], TwoMoons
target, reverse_kld) and of the register-resident MLP kernels alone."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import vcnf_b200 as nf
from vcnf_b200 import ops

dev = "cuda"
torch.manual_seed(0)
flows = []
for _ in range(16):
    flows += [nf.flows.AffineCouplingBlock(nf.nets.MLP([1, 32, 32, 2], init_zeros=False)), nf.flows.Permute(2, mode="swap")]
model = nf.NormalizingFlow(nf.distributions.DiagGaussian(2), flows, nf.distributions.TwoMoons()).to(dev)
opt = torch.optim.Adam(model.parameters(), lr=1e-4)
for B in (1024, 1 << 20):
    def step():
        opt.zero_grad(set_to_none=True)
        loss = model.reverse_kld(num_samples=B)
        loss.backward()
        opt.step()
        return loss
    for _ in range(3):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 10
    launches = ops.LAUNCHES
    e0.record()
    for _ in range(n):
        loss = step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    print("cfg1 reverse_kld B=%d: %.3f ms/step eager, %.2f M samples/s, loss %.4f, %d vcnf launches/step" % (
        B, ms, B / ms / 1e3, float(loss), (ops.LAUNCHES - launches) // n))
# the same step replayed from a CUDA graph (parallel.GraphedTrainStep with a reverse_kld loss); the eager loss must
# be released first: it keeps AccumulateGrad nodes bound to the legacy stream alive, which capture cannot wait on
del loss
from vcnf_b200 import parallel
for B in (1024, 1 << 16, 1 << 20):
    opt_g = torch.optim.Adam(model.parameters(), lr=1e-4, capturable=True)
    bucket = parallel.GradBucket(model.parameters())
    dummy = torch.zeros(1, device=dev)
    gstep = parallel.GraphedTrainStep(model, opt_g, bucket, dummy, world=1,
                                      loss_fn=lambda m, x, B=B: m.reverse_kld(num_samples=B))
    if gstep.graph is None:
        print("capture failed:", gstep.capture_error[:600])
        break
    for _ in range(3):
        gstep(dummy)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50):
        gloss = gstep(dummy)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 50
    print("cfg1 reverse_kld B=%d graphed: %.3f ms/step, %.2f M samples/s, loss %.4f" % (B, ms, B / ms / 1e3, float(gloss)))
    del gstep, gloss
# the MLP kernels alone
B = 1 << 20
net = nf.nets.MLP([1, 32, 32, 2]).to(dev)
x = torch.randn(B, 1, device=dev, requires_grad=True)
gy = torch.randn(B, 2, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for name in ("fwd", "bwd"):
    ts = []
    for _ in range(8):
        y = net(x)
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if name == "fwd":
            with torch.no_grad():
                e0.record(); net(x); e1.record()
        else:
            e0.record(); torch.autograd.grad(y, [x] + list(net.parameters()), gy); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    us = sorted(ts)[len(ts) // 2]
    flop = B * 1120 * 2 * (1 if name == "fwd" else 3)
    print("mlp[1,32,32,2] %s B=2^20: %.1f us  %.1f TFLOP/s fp32 (FMA pipe)  %.0f GB/s algorithmic (%d B/sample)" % (
        name, us, flop / us / 1e6, B * (12 if name == "fwd" else 16) / us / 1e3, 12 if name == "fwd" else 16))
