"""Kernel durations (torch profiler) of the register MLP forward / backward over the batch size, on both sides of
kMlpQuadMaxRows (four lanes per sample below it, one thread per sample above)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from torch.profiler import profile, ProfilerActivity
import vcnf_b200 as nf

dev = torch.device("cuda", 0)
torch.manual_seed(0)
for dims, dt in (([1, 32, 32, 2], torch.float32), ([15, 30, 15], torch.float64)):
    net = nf.nets.MLP(dims).to(dt).to(dev)
    for B in (1024, 2048, 4096, 8192, 16384, 16385, 32768, 65536):
        x = torch.randn(B, dims[0], device=dev, dtype=dt, requires_grad=True)
        for _ in range(3):
            net(x).sum().backward()
        torch.cuda.synchronize()
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            for _ in range(10):
                net(x).sum().backward()
            torch.cuda.synchronize()
        row = []
        for e in prof.key_averages():
            if "mlp_" in e.key:
                row.append("%s %.1f us" % (e.key.split("vcnf::")[1].split("<")[0], e.device_time_total / e.count))
        print(dims, str(dt).split(".")[1], B, " | ".join(sorted(row)), flush=True)
