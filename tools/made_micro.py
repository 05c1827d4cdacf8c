"""One sampling-direction call of a cfg-3 layer (for ncu): MaskedAffineAutoregressive(64, 256) and
AutoregressiveRationalQuadraticSpline(64, 2, 256) at B = 32768."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import vcnf_b200 as nf
torch.manual_seed(0)
maf = nf.flows.MaskedAffineAutoregressive(64, 256).cuda()
ars = nf.flows.AutoregressiveRationalQuadraticSpline(64, 2, 256).cuda().mprqat
z = torch.randn(32768, 64, device="cuda")
with torch.no_grad():
    for layer in (maf, ars):
        layer.inverse(z)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); layer.inverse(z); e1.record(); torch.cuda.synchronize()
        print(type(layer).__name__, "inverse(32768): %.3f ms" % e0.elapsed_time(e1))
