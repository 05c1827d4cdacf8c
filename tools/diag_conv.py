"""Diagnostic: per-tensor error of the bf16 ConvNet2d route against the library convolution in fp64."""
import copy, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import vcnf_b200 as nf
from vcnf_b200.nets import cnn

def rel(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).abs().max() / (b.abs().max() + 1e-30))

for channels, B, H, W, leaky in [((6, 64, 64, 12), 5, 16, 16, 0.0), ((12, 256, 256, 24), 3, 8, 8, 0.0), ((24, 128, 128, 48), 9, 4, 4, 0.1)]:
    torch.manual_seed(0)
    net = nf.nets.ConvNet2d(channels, (3, 1, 3), leaky=leaky, init_zeros=False)
    with torch.no_grad():
        for p in net.parameters():
            p.mul_(1.5)
    ref = copy.deepcopy(net.net).double()
    net = net.cuda()
    g = torch.Generator().manual_seed(9)
    x = torch.randn(B, channels[0], H, W, generator=g)
    gy = torch.randn(B, channels[-1], H, W, generator=g)
    xr = x.double().requires_grad_(True)
    yr = ref(xr); (yr * gy.double()).sum().backward()
    cnn.set_precision("bf16")
    xg = x.cuda().requires_grad_(True)
    y = net(xg); (y * gy.cuda()).sum().backward()
    print(channels, "y", rel(y, yr), "gx", rel(xg.grad, xr.grad))
    for (n, p), (_, q) in zip(net.net.named_parameters(), ref.named_parameters()):
        print("   ", n, rel(p.grad, q.grad))
    e = (xg.grad.cpu().double() - xr.grad).abs()
    print("   gx err by h:", [round(float(v), 3) for v in e.amax(dim=(0, 1, 3))])
    print("   gx err by w:", [round(float(v), 3) for v in e.amax(dim=(0, 1, 2))])
    print("   gx err by c:", [round(float(v), 3) for v in e.amax(dim=(0, 2, 3))])
    print("   gx err by b:", [round(float(v), 3) for v in e.amax(dim=(1, 2, 3))], "scale", float(xr.grad.abs().max()))
