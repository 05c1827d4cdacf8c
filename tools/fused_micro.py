"""Times one cfg-2 spline coupling layer in the no-grad directions: the fused kernel (vcnf_resnet_spline_fwd, with and
without its pack launch) against the layer-by-layer kernels, CUDA events."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import vcnf_b200 as nf
from vcnf_b200 import ops

dev = torch.device("cuda", 0)
torch.manual_seed(0)
layer = nf.flows.CoupledRationalQuadraticSpline(29, 2, 128).to(dev)
c = layer.prqct
B = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 20
x = torch.randn(B, 29, device=dev)


def timed(fn, n=iters):
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


with torch.no_grad():
    fused = c._fused_route(x, None)
    w, b = c.transform_net.fused_params()
    shared = c._shared_params()
    for inverse in (False, True):
        t_all = timed(lambda: ops.resnet_spline(x, fused, w, b, inverse=inverse, shared=shared))
        packed = fused.pack(w, b)
        fused.pack = lambda *_: packed
        t_k = timed(lambda: ops.resnet_spline(x, fused, w, b, inverse=inverse, shared=shared))
        del fused.pack
        c._fused = False
        t_ref = timed(lambda: (c.inverse if inverse else c.forward)(x))
        c._fused = fused
        print("B=%d inverse=%d  fused %.1f us (kernel alone %.1f us)  layer-by-layer %.1f us" % (B, inverse, t_all, t_k, t_ref),
              flush=True)
