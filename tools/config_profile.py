"""GPU-box diagnostic: device time per kernel of one EAGER training step of a bench configuration (torch profiler,
warm caches), e.g. `python tools/config_profile.py cfg5 [batch]`.  The graphed step replays the same kernels."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
import vcnf_b200 as nf
from vcnf_b200 import parallel

name = sys.argv[1]
dev = torch.device("cuda", 0)
spec = bench.config_spec(name, nf)
if spec.get("setup"):
    spec["setup"]()
B = int(sys.argv[2]) if len(sys.argv) > 2 else spec["batch"]
torch.manual_seed(0)
model = spec["build"]().to(dev)
x = spec["data"](B, 1234).to(dev)
loss_fn = spec["loss_fn"] or (lambda m, xb: m.forward_kld(xb))
if name == "cfg1":
    loss_fn = lambda m, xb, B=B: m.reverse_kld(num_samples=B)          # noqa: E731
if loss_fn == "xy":
    y = torch.randint(0, 10, (B,), generator=torch.Generator().manual_seed(99)).to(dev)
    loss_fn = lambda m, xb, y=y: m.forward_kld(xb, y)                  # noqa: E731
with torch.no_grad():
    loss_fn(model, x.clone())
if spec.get("optimizer") == "adamax":
    opt = nf.optim.FlatAdamax(model.parameters(), lr=spec["lr"], weight_decay=spec["wd"])
    bucket = opt.bucket
else:
    opt = nf.optim.FlatAdam(model.parameters(), lr=spec["lr"], weight_decay=spec["wd"])
    bucket = opt.bucket


def step():
    bucket.begin()
    loss = loss_fn(model, x.clone())
    loss.backward()
    opt.step()
    return loss


for _ in range(3):
    step()
torch.cuda.synchronize()
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    for _ in range(2):
        step()
    torch.cuda.synchronize()
ev = [e for e in prof.key_averages() if e.self_device_time_total > 0 and e.cpu_time_total == 0]   # kernel rows only
tot = sum(e.self_device_time_total for e in ev)
print("# %s, batch %d, eager step: device busy %.2f ms over %d kernel launches per step (torch profiler, warm)"
      % (name, B, tot / 2e3, sum(e.count for e in ev) / 2))
print("#   ms/step  share  launches/step  us/launch  kernel")
for e in sorted(ev, key=lambda e: -e.self_device_time_total)[:60]:
    print("%9.3f %5.1f%%  %6d  %8.1f  %s" % (e.self_device_time_total / 2e3, 100 * e.self_device_time_total / tot, e.count / 2,
                                         e.self_device_time_total / e.count, e.key[:110]))
