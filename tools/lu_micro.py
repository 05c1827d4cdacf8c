"""LULinearPermute(29) density direction + backward at B = 65536 (cfg 2 layer), for ncu."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import vcnf_b200 as nf
torch.manual_seed(0)
layer = nf.flows.LULinearPermute(29, identity_init=False).cuda()
z = torch.randn(65536, 29, device="cuda", requires_grad=True)
for _ in range(3):
    y, ld = layer.inverse(z)
    (y.sum() + ld.sum()).backward()
torch.cuda.synchronize()
print("done")
