"""GPU-box diagnostic: device time by kernel of one cfg-2 sample(65536) call (sampling direction)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
import vcnf_b200 as nf
from torch.profiler import profile, ProfilerActivity
dev = torch.device("cuda", 0)
torch.manual_seed(0)
model = bench.build_model(nf).to(dev)
with torch.no_grad():
    for _ in range(3):
        model.sample(bench.BATCH)
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
        model.sample(bench.BATCH)
        torch.cuda.synchronize()
ev = [e for e in prof.key_averages() if e.self_device_time_total > 0 and e.cpu_time_total == 0]
tot = sum(e.self_device_time_total for e in ev)
print("device busy per draw %.2f ms over %d kernel launches" % (tot / 1e3, sum(e.count for e in ev)))
for e in sorted(ev, key=lambda e: -e.self_device_time_total)[:30]:
    print("%8.3f ms %5.1f%% n=%4d  %s" % (e.self_device_time_total / 1e3, 100 * e.self_device_time_total / tot, e.count, e.key[:100]))
