"""GPU-box timing of the spline layer kernels alone at the bench shapes (CUDA events, L2 flushed)."""
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
cond = ops._padded_empty(B, spec.Tc * spec.Pc, torch.float32, "cuda")
cond.copy_(0.5 * torch.randn(B, spec.Tc * spec.Pc, device="cuda"))
cond.requires_grad_(True)
shared = layer.unconditional_transform.packed_params().detach().clone().requires_grad_(True)
gy, gld = torch.randn(B, D, device="cuda"), torch.randn(B, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
tf, tb = [], []
for i in range(12):
    flush.zero_()
    e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    e[0].record()
    y, ld = ops.rqs_layer(x, cond, shared, spec, False)
    e[1].record()
    flush.zero_()
    e2 = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    e2[0].record()
    torch.autograd.grad([y, ld], [x, cond, shared], [gy, gld])
    e2[1].record()
    torch.cuda.synchronize()
    if i >= 3:
        tf.append(e[0].elapsed_time(e[1]) * 1e3); tb.append(e2[0].elapsed_time(e2[1]) * 1e3)
print("stages=%s fwd %.1f us (min %.1f)  bwd(+reduce, + autograd glue) %.1f us (min %.1f)" % (
    os.environ.get("VCNF_RQS_STAGES", "2"), sum(tf) / len(tf), min(tf), sum(tb) / len(tb), min(tb)))
