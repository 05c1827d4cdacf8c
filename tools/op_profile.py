"""GPU-box diagnostic: ATen-level view of one eager cfg-2 training step (which torch ops still launch kernels
around the hand-written ones), sorted by device time, with input shapes."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
import vcnf_b200 as nf
from vcnf_b200 import parallel
from torch.profiler import profile, ProfilerActivity

dev = torch.device("cuda", 0)
torch.manual_seed(0)
model = bench.build_model(nf).to(dev)
bucket = parallel.GradBucket(model.parameters())
opt = torch.optim.Adam(model.parameters(), lr=1e-3, fused=True)
x = bench.synth_batch(bench.BATCH, 1234).to(dev)

def step():
    bucket.begin()
    loss = model.forward_kld(x)
    loss.backward()
    opt.step()

for _ in range(3):
    step()
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], record_shapes=True) as prof:
    step()
    torch.cuda.synchronize()
rows = [e for e in prof.key_averages(group_by_input_shape=True) if e.self_device_time_total > 0 and e.cpu_time_total > 0]
for e in sorted(rows, key=lambda e: -e.self_device_time_total)[:40]:
    print("%8.3f ms n=%4d  %-40s %s" % (e.self_device_time_total / 1e3, e.count, e.key[:40], str(e.input_shapes)[:110]))
