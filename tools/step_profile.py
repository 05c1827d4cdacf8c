"""GPU-box diagnostic: where one cfg-2 training step spends device time vs host wall time."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
import vcnf_b200 as nf
from vcnf_b200 import parallel

B = int(sys.argv[1]) if len(sys.argv) > 1 else bench.BATCH
dev = torch.device("cuda", 0)
torch.manual_seed(0)
model = bench.build_model(nf).to(dev)
bucket = parallel.GradBucket(model.parameters())
opt = torch.optim.Adam(model.parameters(), lr=1e-3, fused=True)
x = bench.synth_batch(B, 1234).to(dev)

def step():
    bucket.begin()
    loss = model.forward_kld(x)
    loss.backward()
    opt.step()
    return loss

for _ in range(3):
    step()
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(5):
    step()
torch.cuda.synchronize()
print("B=%d wall per step %.2f ms" % (B, (time.perf_counter() - t0) / 5 * 1e3))
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    for _ in range(2):
        step()
    torch.cuda.synchronize()
ev = [e for e in prof.key_averages() if e.self_device_time_total > 0 and e.cpu_time_total == 0]   # kernel rows only
tot = sum(e.self_device_time_total for e in ev)
print("device busy per step %.2f ms over %d kernel launches per step" % (tot / 2e3, sum(e.count for e in ev) / 2))
for e in sorted(ev, key=lambda e: -e.self_device_time_total)[:45]:
    print("%8.3f ms %5.1f%% n=%4d  %s" % (e.self_device_time_total / 2e3, 100 * e.self_device_time_total / tot, e.count / 2, e.key[:90]))
