"""One launch of each non-GEMM / non-spline kernel at its bench shape (for `ncu --set full`): the affine family,
per-row linear maps, optimiser, base distributions, dequantiser, convolution kernels.  Run under ncu with
--profile-from-start off; the profiled region is bracketed by cudaProfilerStart/Stop."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests", "golden")]
import torch
import bench
from vcnf_b200 import ops
import vcnf_b200 as nf

dev = torch.device("cuda", 0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
fns = []


def grab(fn, flush, reps=20):
    fn()                               # warm (module load, attribute setting) outside the profiled range
    torch.cuda.synchronize()
    flush.zero_()
    torch.cuda.cudart().cudaProfilerStart()
    fn()
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStop()
    return 1.0


bench._time_launches = grab
bench.extra_kernel_rooflines(dev, flush)
bench.conv_roofline(dev, flush)
# dequantiser (cfg 4 shape: 10 columns, 24 layers, B = 65536) and the base distributions
B, n, L = 65536, 10, 24
x = torch.randint(0, 5, (B, n), device=dev).float()
noise = torch.rand(B, n, dtype=torch.float64, device=dev)
tab = (0.1 * torch.randn(n, L, 2, dtype=torch.float64, device=dev)).requires_grad_(True)
q = torch.full((n,), 5.0, dtype=torch.float64, device=dev)


def vdeq():
    zc, lq = ops.vdeq_fused(x, noise, tab, q, 1e-5)
    (zc.sum() + lq.sum()).backward()


gm = nf.distributions.GaussianMixture(100, 15).double().to(dev)
z15 = torch.randn(B, 15, dtype=torch.float64, device=dev, requires_grad=True)
dg = nf.distributions.DiagGaussian(29).to(dev)
z29 = torch.randn(B, 29, device=dev, requires_grad=True)
fns += [vdeq, lambda: gm.log_prob(z15).sum().backward(), lambda: dg.log_prob(z29).sum().backward()]
for f in fns:
    grab(f, flush)
print("done")
