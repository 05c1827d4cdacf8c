"""Cost of the training outputs of the fused net + spline kernel: none / layer outputs / parameters / both."""
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
x = torch.randn(B, 29, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
with torch.no_grad():
    fused = c._fused_route(x, None)
    w, b = c.transform_net.fused_params()
    shared = c._shared_params()
    packed = fused.pack(w, b)
    fused.pack = lambda *_: packed
    for wp, wa in ((False, False), (False, True), ("padded", False), ("padded", True)):
        fn = lambda: ops.resnet_spline(x, fused, w, b, inverse=False, shared=shared, want_params=wp, want_acts=wa)
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        tot = 0.0
        for _ in range(10):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record()
            torch.cuda.synchronize()
            tot += e0.elapsed_time(e1)
        print("params=%s acts=%s: %.1f us" % (wp, wa, tot / 10 * 1e3), flush=True)
