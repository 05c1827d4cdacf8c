"""GPU-box diagnostic: idle time between consecutive kernels of a REPLAYED training step (torch profiler timeline of
graph replays): which kernel pairs leave the GPU idle, and for how long.  usage: python tools/graph_gaps.py [cfg2|cfg3a|...]"""
import os, sys, collections
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
import vcnf_b200 as nf
from vcnf_b200 import parallel

name = sys.argv[1] if len(sys.argv) > 1 else "cfg2"
dev = torch.device("cuda", 0)
torch.manual_seed(0)
if name == "cfg2":
    model = bench.build_model(nf).to(dev)
    x = bench.synth_batch(bench.BATCH, 1234).to(dev)
    opt = nf.optim.FlatAdam(model.parameters(), lr=1e-3)
    loss_fn = None
else:
    spec = bench.config_spec(name, nf)
    if spec.get("setup"):
        spec["setup"]()
    model = spec["build"]().to(dev)
    x = spec["data"](spec["batch"], 1234).to(dev)
    loss_fn = spec["loss_fn"]
    if loss_fn == "xy":
        y = torch.randint(0, 10, (spec["batch"],), generator=torch.Generator().manual_seed(99)).to(dev)
        loss_fn = lambda m, xb, y=y: m.forward_kld(xb, y)              # noqa: E731
    if name == "cfg1":
        loss_fn = lambda m, xb: m.reverse_kld(num_samples=spec["batch"])   # noqa: E731
    with torch.no_grad():
        (loss_fn or (lambda m, xb: m.forward_kld(xb)))(model, x.clone())
    opt = (nf.optim.FlatAdamax if spec.get("optimizer") == "adamax" else nf.optim.FlatAdam)(
        model.parameters(), lr=spec["lr"], weight_decay=spec["wd"])
step = parallel.GraphedTrainStep(model, opt, opt.bucket, x, world=1, loss_fn=loss_fn)
assert step.graph is not None, step.capture_error
for _ in range(3):
    step(x)
torch.cuda.synchronize()
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(3):
        step(x)
    torch.cuda.synchronize()
ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA and e.time_range.end > e.time_range.start]
ev.sort(key=lambda e: e.time_range.start)
# last replay only
n = len(ev) // 3
ev = ev[-n:]
busy = sum(e.time_range.end - e.time_range.start for e in ev)
span = ev[-1].time_range.end - ev[0].time_range.start
gaps = collections.defaultdict(lambda: [0, 0.0])
overlap = 0.0
for a, b in zip(ev, ev[1:]):
    g = b.time_range.start - a.time_range.end
    if g < 0:
        overlap -= g
        continue
    k = (a.name[:60], b.name[:60])
    gaps[k][0] += 1
    gaps[k][1] += g
tot_gap = sum(v[1] for v in gaps.values())
print("# %s: one replayed step: %d kernels, span %.3f ms, sum of kernel durations %.3f ms, idle gaps %.3f ms, overlap (dependent "
      "launch) %.3f ms" % (name, len(ev), span / 1e3, busy / 1e3, tot_gap / 1e3, overlap / 1e3))
print("# idle us  count  us/gap   after kernel  ->  before kernel")
for k, (c, t) in sorted(gaps.items(), key=lambda kv: -kv[1][1])[:40]:
    print("%8.1f %5d %7.2f   %s  ->  %s" % (t, c, t / c, k[0], k[1]))
agg = collections.defaultdict(lambda: [0, 0.0])
for e in ev:
    agg[e.name[:110]][0] += 1
    agg[e.name[:110]][1] += e.time_range.end - e.time_range.start
print("# kernel durations inside the replayed step (a dependent-launch kernel's duration includes its wait)")
print("#   ms/step  share  launches  us/launch  kernel")
for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:45]:
    print("%9.3f %5.1f%%  %6d  %8.1f  %s" % (t / 1e3, 100 * t / busy, c, t / c, k))
if len(sys.argv) > 2:
    lo, hi = int(sys.argv[2]), int(sys.argv[3])
    print("# kernel sequence [%d, %d) of the replayed step: start us (relative), duration us, name" % (lo, hi))
    t0 = ev[lo].time_range.start
    for e in ev[lo:hi]:
        print("%9.1f %7.1f  %s" % (e.time_range.start - t0, e.time_range.end - e.time_range.start, e.name[:100]))
