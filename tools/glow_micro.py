"""The small kernels of a cfg-5 GlowBlock at the three level shapes (for ncu / CUDA-event timing): Invertible1x1Conv in the
density direction with its backward (glow_w, small_linear, small_linear_wgrad_img)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import vcnf_b200 as nf

B = int(sys.argv[1]) if len(sys.argv) > 1 else 128
dev = torch.device("cuda", 0)
torch.manual_seed(0)
for C, S in ((48, 4), (24, 8), (12, 16)):
    layer = nf.flows.Invertible1x1Conv(C, use_lu=True).to(dev)
    z = torch.randn(B, C, S, S, device=dev, requires_grad=True)

    def step():
        y, ld = layer.inverse(z)
        (y.sum() + ld).backward()
    for _ in range(3):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        step()
    e1.record()
    torch.cuda.synchronize()
    print("C=%d %dx%d B=%d: fwd+bwd %.1f us" % (C, S, S, B, 1e3 * e0.elapsed_time(e1) / 20))
