"""GPU-box timing of cfg 5 (example/glow.ipynb:37-73: L = 3, K = 16, hidden 256, ClassCondDiagGaussian bases, 10 classes,
3 x 32 x 32 inputs, MultiscaleFlow.forward_kld(x, y), Adamax(1e-3, weight_decay 1e-5)).  The bijector math (coupling,
ActNorm, invertible 1x1 / LU, squeeze, split / merge) runs on this package's kernels; the 3x3 / 1x1 convolutions of
ConvNet2d still run on cuDNN (partial)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import torch
import vcnf_b200 as nf
from vcnf_b200 import ops
from cases import build_glow_small

dev = "cuda"
torch.manual_seed(0)
from vcnf_b200.nets import cnn
cnn.set_precision(os.environ.get("VCNF_CONV_PRECISION", "bf16"))
print("ConvNet2d precision:", cnn.get_precision())
model = build_glow_small(nf, L=3, K=16, hidden=256, in_shape=(3, 32, 32), classes=10).to(dev)
opt = torch.optim.Adamax(model.parameters(), lr=1e-3, weight_decay=1e-5)
for B in (128, 512):
    x = (torch.rand(B, 3, 32, 32) * 255 / 256 + torch.rand(B, 3, 32, 32) / 256).to(dev)
    y = torch.randint(0, 10, (B,)).to(dev)

    def step():
        opt.zero_grad(set_to_none=True)
        loss = model.forward_kld(x, y)
        loss.backward()
        opt.step()
        return loss
    for _ in range(4):                       # the first call runs ActNorm's data-dependent initialisation
        loss = step()
    torch.cuda.synchronize()
    l0 = ops.LAUNCHES
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 8
    e0.record()
    for _ in range(n):
        loss = step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    print("cfg5 Glow forward_kld train step B=%d (eager): %.2f ms/step, %.0f images/s, loss %.3f, %d vcnf launches/step" % (
        B, ms, B / ms * 1e3, float(loss), (ops.LAUNCHES - l0) // n))
# the same step replayed from a CUDA graph
from vcnf_b200 import parallel
del loss
for B in (128, 512):
    x = (torch.rand(B, 3, 32, 32) * 255 / 256 + torch.rand(B, 3, 32, 32) / 256).to(dev)
    y = torch.randint(0, 10, (B,)).to(dev)
    opt_g = torch.optim.Adamax(model.parameters(), lr=1e-3, weight_decay=1e-5, capturable=True)
    bucket = parallel.GradBucket(model.parameters())
    gstep = parallel.GraphedTrainStep(model, opt_g, bucket, x, world=1, loss_fn=lambda m, xb, y=y: m.forward_kld(xb, y))
    if gstep.graph is None:
        print("cfg5 capture failed:", (gstep.capture_error or "")[:400])
        break
    for _ in range(3):
        gstep(x)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(8):
        gl = gstep(x)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 8
    print("cfg5 Glow forward_kld train step B=%d (graph): %.2f ms/step, %.0f images/s, loss %.3f" % (B, ms, B / ms * 1e3, float(gl)))
    del gstep, gl
