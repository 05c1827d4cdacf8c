"""Kernel durations inside one replay of the graphed cfg-2 sampler (torch profiler)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from torch.profiler import profile, ProfilerActivity
import vcnf_b200 as nf
from vcnf_b200 import parallel
import bench

dev = torch.device("cuda", 0)
torch.manual_seed(0)
model = bench.build_model(nf).to(dev)
draw = parallel.GraphedSampler(model, bench.BATCH)
for _ in range(3):
    draw()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    draw()
    torch.cuda.synchronize()
ev = sorted([e for e in prof.events() if e.device_time > 0], key=lambda e: e.time_range.start)
span = ev[-1].time_range.end - ev[0].time_range.start
print("# one replay: %d kernels, span %.3f ms" % (len(ev), span / 1e3))
agg = {}
for e in ev:
    a = agg.setdefault(e.name[:110], [0, 0.0]); a[0] += 1; a[1] += e.device_time
for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:18]:
    print("%8.3f ms %5d x %7.1f us  %s" % (t / 1e3, c, t / c, k))
