"""GPU-box timing of cfg 4 (runadultvdeq1ft.py recipe: d = 15, 10 categorical columns each with a 1-D variational
dequantiser of 8 x [ActNorm, MaskedAffineFlow(MLP[1,4,1]), ActNorm], main flow 12 x [MaskedAffineFlow(MLP[15,30,15] x 2),
ActNorm(15)], GaussianMixture base, fp64 model, fp32 input, forward_kld + Adam(1e-4, weight_decay 1e-6))."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import torch
import vcnf_b200 as nf
from vcnf_b200 import ops
from cases import build_cfg4_small, cfg4_data

dev = "cuda"
torch.manual_seed(0)
cats = (1, 3, 4, 5, 6, 7, 8, 9, 13, 14)
levels = (9, 16, 16, 7, 15, 6, 5, 2, 42, 2)
model = build_cfg4_small(nf, d=15, cats=cats, levels=levels, n_main=12, n_vdeq=8).double().to(dev)
opt = torch.optim.Adam(model.parameters(), lr=1e-4, weight_decay=1e-6)
for B in (2048, 65536):
    x0 = cfg4_data(n=B, d=15, cats=cats, levels=levels).to(dev)

    def step():
        opt.zero_grad(set_to_none=True)
        loss = model.forward_kld(x0.clone())          # forward_kld dequantises its argument in place (core.py:54)
        loss.backward()
        opt.step()
        return loss
    for _ in range(4):
        loss = step()
    torch.cuda.synchronize()
    l0 = ops.LAUNCHES
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 5
    e0.record()
    for _ in range(n):
        loss = step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    print("cfg4 forward_kld train step B=%d fp64 (eager): %.2f ms/step, %.3f M samples/s, loss %.4f, %d vcnf launches/step" % (
        B, ms, B / ms / 1e3, float(loss), (ops.LAUNCHES - l0) // n))
# the same step replayed from a CUDA graph
from vcnf_b200 import parallel
del loss
for B in (2048, 65536):
    x0 = cfg4_data(n=B, d=15, cats=cats, levels=levels).to(dev)
    opt_g = nf.optim.FlatAdam(model.parameters(), lr=1e-4, weight_decay=1e-6)
    gstep = parallel.GraphedTrainStep(model, opt_g, opt_g.bucket, x0)
    if gstep.graph is None:
        print("cfg4 capture failed:", (gstep.capture_error or "")[:400])
        break
    for _ in range(3):
        gstep(x0)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        gl = gstep(x0)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print("cfg4 forward_kld train step B=%d fp64 (graph): %.2f ms/step, %.3f M samples/s, loss %.4f" % (B, ms, B / ms / 1e3, float(gl)))
    del gstep, gl
