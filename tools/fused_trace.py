"""Timeline of one tile of the fused net + spline kernel.  Needs a library built with the trace points compiled in:
    touch vcnf_b200/csrc/resnet_fused.cu && VCNF_NVCC_EXTRA=-DVCNF_FUSED_TRACE python -m vcnf_b200.build
(rebuild without the variable afterwards: the trace stores cost ~7 % of the kernel)."""
import os, sys, ctypes
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import vcnf_b200 as nf
from vcnf_b200 import ops, _lib
dev = torch.device("cuda", 0)
torch.manual_seed(0)
layer = nf.flows.CoupledRationalQuadraticSpline(29, 2, 128).to(dev)
c = layer.prqct
B = int(sys.argv[1]) if len(sys.argv) > 1 else 18944
x = torch.randn(B, 29, device=dev)
with torch.no_grad():
    for _ in range(3):
        c.forward(x)
    torch.cuda.synchronize()
buf = (ctypes.c_longlong * 256)()
lib = _lib.lib()
lib.vcnf_debug_fused_trace.argtypes = [ctypes.c_void_p]
print("rc", lib.vcnf_debug_fused_trace(buf))
mma = [buf[i] for i in range(128)]
epi = [buf[128 + i] for i in range(128)]
t0 = min(v for v in mma + epi if v > 0)
print("MMA  :", [round((v - t0) / 1.9e3, 2) if v else None for v in mma[:80]])
print("EPI  :", [round((v - t0) / 1.9e3, 2) if v else None for v in epi[:60]])
