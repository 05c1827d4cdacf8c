"""GPU-box timing of cfg 3a (8 x [AutoregressiveRationalQuadraticSpline(64, 2, 256), LULinearPermute(64)],
DiagGaussian(64, trainable=False), x = 1.5 randn(B, 64), B = 32 768 per GPU): forward_kld training step replayed from a
CUDA graph, and sample() (64 conditioner passes per layer)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import vcnf_b200 as nf
from vcnf_b200 import ops, parallel

dev = "cuda"
torch.manual_seed(0)
B, D = 32768, 64
flows = []
for _ in range(8):
    flows += [nf.flows.AutoregressiveRationalQuadraticSpline(D, 2, 256), nf.flows.LULinearPermute(D)]
model = nf.NormalizingFlow(nf.distributions.DiagGaussian(D, trainable=False), flows).to(dev)
x = (1.5 * torch.randn(B, D)).to(dev)
opt = nf.optim.FlatAdam(model.parameters(), lr=1e-3)
step = parallel.GraphedTrainStep(model, opt, opt.bucket, x)
print("graph:", step.graph is not None, step.capture_error)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
for _ in range(3):
    step(x)
torch.cuda.synchronize()
l0 = ops.LAUNCHES
step._eager()
per_step = ops.LAUNCHES - l0
ts = []
for _ in range(10):
    flush.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); loss = step(x); e1.record()
    torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
ms = sum(ts) / len(ts)
print("cfg3a forward_kld train step B=%d: %.3f ms/step, %.2f M samples/s, loss %.4f, %d vcnf launches/step" % (
    B, ms, B / ms / 1e3, float(loss), per_step))
with torch.no_grad():
    for n in (4096, 32768):
        model.sample(n)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); z, lq = model.sample(n); e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        print("cfg3a sample(%d): %.1f ms, %.3f M samples/s, finite %s" % (n, ms, n / ms / 1e3, bool(torch.isfinite(z).all())))
draw = parallel.GraphedSampler(model, 32768)
print("sampler graph:", draw.graph is not None, draw.capture_error)
for _ in range(2):
    draw()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); z, lq = draw(); e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1)
print("cfg3a sample(32768) replayed from a graph: %.1f ms, %.3f M samples/s, finite %s" % (
    ms, 32768 / ms / 1e3, bool(torch.isfinite(z).all())))
# cfg 3b: 8 x [Reverse(MaskedAffineAutoregressive(64, 256)), Permute(64, 'shuffle')]
del step, draw
torch.manual_seed(0)
flows = []
for _ in range(8):
    flows += [nf.flows.Reverse(nf.flows.MaskedAffineAutoregressive(D, 256)), nf.flows.Permute(D, mode="shuffle")]
model_b = nf.NormalizingFlow(nf.distributions.DiagGaussian(D, trainable=False), flows).to(dev)
opt_b = nf.optim.FlatAdam(model_b.parameters(), lr=1e-3)
step_b = parallel.GraphedTrainStep(model_b, opt_b, opt_b.bucket, x)
print("cfg3b graph:", step_b.graph is not None, step_b.capture_error)
for _ in range(3):
    step_b(x)
torch.cuda.synchronize()
ts = []
for _ in range(10):
    flush.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); loss = step_b(x); e1.record()
    torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
ms = sum(ts) / len(ts)
print("cfg3b forward_kld train step B=%d: %.3f ms/step, %.2f M samples/s, loss %.4f" % (B, ms, B / ms / 1e3, float(loss)))
draw_b = parallel.GraphedSampler(model_b, 32768)
for _ in range(2):
    draw_b()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); z, lq = draw_b(); e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1)
print("cfg3b sample(32768) graph=%s: %.1f ms, %.3f M samples/s, finite %s" % (
    draw_b.graph is not None, ms, 32768 / ms / 1e3, bool(torch.isfinite(z).all())))
