"""Short driver for ncu: the LU parameter kernels (compose fwd / bwd, inverse) at D = 29."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import vcnf_b200 as nf
torch.manual_seed(0)
lu = nf.flows.LULinearPermute(29, identity_init=False).cuda()
x = torch.randn(4096, 29, device="cuda")
for _ in range(3):
    with torch.no_grad():
        lu.forward(x)
    z, ld = lu.inverse(x)
    (z.sum() + ld.sum()).backward()
torch.cuda.synchronize()
print("ok")
