"""cfg 2 as the run scripts build it: `nfm.double()` (runcred.py:151) -- the step whose conditioners leave the tcgen05 path
(no fp64-faithful tensor-core mode; cuBLAS DGEMM under the same modules).  Graph replay, B = 65536 and 8192."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import bench
import vcnf_b200 as nf
from vcnf_b200 import parallel

dev = torch.device("cuda", 0)
for B in (8192, 65536):
    torch.manual_seed(0)
    model = bench.build_model(nf).double().to(dev)
    x = bench.synth_batch(B, 1234).double().to(dev)
    opt = nf.optim.FlatAdam(model.parameters(), lr=1e-3)
    step = parallel.GraphedTrainStep(model, opt, opt.bucket, x, world=1)
    for _ in range(3):
        step(x)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5):
        loss = step(x)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 5
    print("cfg2 fp64, B = %d: %.2f ms/step = %.0f samples/s (graph: %s), loss %.6f" % (B, ms, B / ms * 1e3, step.graph is not None, float(loss)))
