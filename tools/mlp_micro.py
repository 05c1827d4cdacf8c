"""One forward + backward of the register-resident MLP [1, 32, 32, 2] on 1 024 rows (cfg 1) and [15, 30, 15] fp64 on 2 048
rows (cfg 4), for ncu."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import vcnf_b200 as nf

dev = torch.device("cuda", 0)
torch.manual_seed(0)
for dims, B, dt in (([1, 32, 32, 2], 1024, torch.float32), ([15, 30, 15], 2048, torch.float64)):
    net = nf.nets.MLP(dims).to(dt).to(dev)
    x = torch.randn(B, dims[0], device=dev, dtype=dt, requires_grad=True)
    for _ in range(3):
        net(x).sum().backward()
    torch.cuda.synchronize()
print("done")
