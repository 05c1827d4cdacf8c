"""Short driver for ncu: the spline layer kernels at the bench shapes (cfg 2: B=65536, D=29, K=8)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import vcnf_b200 as nf
from vcnf_b200 import ops

B, D = 65536, 29
torch.manual_seed(0)
layer = nf.flows.CoupledRationalQuadraticSpline(D, 2, 128).cuda().prqct
layer._build_specs()
spec = layer._specs["both"]
x = (1.5 * torch.randn(B, D, device="cuda")).requires_grad_(True)
cond = (0.5 * torch.randn(B, spec.Tc * spec.Pc, device="cuda")).requires_grad_(True)
shared = layer.unconditional_transform.packed_params().detach().clone().requires_grad_(True)
gy, gld = torch.randn(B, D, device="cuda"), torch.randn(B, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 3):
    flush.zero_()
    y, ld = ops.rqs_layer(x, cond, shared, spec, False)
    torch.autograd.grad([y, ld], [x, cond, shared], [gy, gld])
torch.cuda.synchronize()
print("ok", float(y.sum()))
