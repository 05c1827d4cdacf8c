"""One cfg-2 coupling layer, density direction WITH gradients: forward (fused values + graph) and forward + backward
against the layer-by-layer path, CUDA events."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import vcnf_b200 as nf
from vcnf_b200 import ops

dev = torch.device("cuda", 0)
torch.manual_seed(0)
layer = nf.flows.CoupledRationalQuadraticSpline(29, 2, 128).to(dev)
B = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
x = torch.randn(B, 29, device=dev)


def timed(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


def fwd():
    return layer.inverse(x)


def fwd_bwd():
    z, ld = layer.inverse(x)
    (z.sum() + ld.sum()).backward()


for flag in (True, False):
    ops.FUSED_NET = flag
    print("fused" if flag else "layerwise", "fwd %.1f us  fwd+bwd %.1f us" % (timed(fwd), timed(fwd_bwd)), flush=True)
