"""Short driver for ncu: LU-linear backward (vcnf_small_linear_bwd) at the bench shape (B=65536, D=29)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from vcnf_b200 import ops
B, D = 65536, 29
torch.manual_seed(0)
x = torch.randn(B, D, device="cuda", requires_grad=True)
W = torch.randn(D, D, device="cuda", requires_grad=True)
b = torch.randn(D, device="cuda", requires_grad=True)
perm = torch.randperm(D, device="cuda").int()
gy = torch.randn(B, D, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for _ in range(3):
    flush.zero_()
    y = ops.small_linear(x, W, b, in_idx=perm)
    torch.autograd.grad(y, [x, W, b], gy)
torch.cuda.synchronize()
print("ok")
