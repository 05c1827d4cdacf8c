#!/usr/bin/env python
"""Benchmark of the bijector hot path (BASELINE.json metric: forward_kld train samples/s).

Workload (BASELINE.json configs[1], SURVEY.md 8d cfg 2): Neural Spline Flow,
16 x [CoupledRationalQuadraticSpline(29, 2 blocks, 128 hidden, 8 bins, linear tails, bound 3) +
LULinearPermute(29)], base DiagGaussian(29), batch 65 536 per GPU, fp32, Adam(1e-3); one "step"
= forward_kld + backward + (gradient all-reduce for N > 1) + optimizer step on one synthetic batch
x = 1.5 * randn(B, 29).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]

Prints ONE JSON line (rank 0).  `--impl reference` times the CPU oracle port of the reference
algorithm (oracle/, torch CPU, all host threads) on a bounded sample of the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for p in (ROOT, os.path.join(ROOT, "tests", "golden")):
    if p not in sys.path:
        sys.path.insert(0, p)

import torch  # noqa: E402

D, LAYERS, BLOCKS, HIDDEN, BINS = 29, 16, 2, 128, 8
BATCH = 65536            # per GPU (weak scaling)
CPU_SAMPLE = 4096        # rows of the same batch the CPU arm processes per step
WORKLOAD = "cfg2 NSF 16x[CoupledRQS(29,2,128,K=8)+LULinearPermute(29)] forward_kld train, batch 65536/GPU, fp32"


def build_model(nf):
    flows = []
    for _ in range(LAYERS):
        flows += [nf.flows.CoupledRationalQuadraticSpline(D, BLOCKS, HIDDEN, num_bins=BINS),
                  nf.flows.LULinearPermute(D)]
    q0 = nf.distributions.DiagGaussian(D, trainable=False)
    return nf.NormalizingFlow(q0, flows, categoricals=[], catlevels=[], catvdeqs=[])


def synth_batch(n, seed):
    g = torch.Generator().manual_seed(seed)
    return 1.5 * torch.randn(n, D, generator=g)


# --------------------------------------------------------------------------------- CPU arm
def cpu_reference_run(steps, warmup, sample=CPU_SAMPLE):
    """The reference algorithm's CPU path: oracle port, all host threads."""
    import oracle
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    torch.manual_seed(0)
    model = build_model(oracle)
    opt = torch.optim.Adam(model.parameters(), lr=1e-3)
    x = synth_batch(sample, 1234)
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        opt.zero_grad(set_to_none=True)
        loss = model.forward_kld(x.clone())
        loss.backward()
        opt.step()
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    total = sum(times)
    return {"value": sample * len(times) / total, "unit": "samples/s", "cores": cores, "kind": "port",
            "sample": "%d of the %d-row batch per step, %d timed steps after %d warm-up, torch CPU %d threads"
                      % (sample, BATCH, len(times), warmup, torch.get_num_threads()),
            "ms_per_step": 1e3 * total / len(times), "loss": float(loss.detach())}


# --------------------------------------------------------------------------------- clocks
class ClockSampler(threading.Thread):
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                f = [s.strip() for s in out.strip().split(",")]
                if len(f) >= 6:
                    self.rows.append(f)
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        self.stop_flag = True
        self.join(timeout=6)
        if not self.rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        mhz = sorted(float(r[0]) for r in self.rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(r[2 + k].lower().startswith("active") for r in self.rows)]
        return {"sm_mhz": mhz[len(mhz) // 2], "sm_max_mhz": float(self.rows[0][1]), "reasons": reasons,
                "samples": len(self.rows)}


# --------------------------------------------------------------------------------- GPU arm
def _peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def _time_launches(fn, flush, reps=20):
    """Average device time of fn() in ms: CUDA events on the launching stream, L2 flushed between."""
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    tot = 0.0
    for _ in range(reps):
        flush.zero_()      # L2 flush; two passes also keep the GPU busy while the host enqueues
        flush.zero_()      # e0 / fn / e1, so the event pair brackets the kernel(s) without a launch gap
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / reps


def kernel_roofline(model, dev, flush):
    """Roofline of the dominant hand-written kernels at the bench shapes, timed live."""
    from vcnf_b200 import ops
    layer = model.flows[0].prqct
    if layer._specs is None:
        layer._build_specs()
    spec = layer._specs["both"]
    B = BATCH
    from vcnf_b200 import _lib
    from vcnf_b200._lib import ptr
    P = spec.Tc * spec.Pc
    x = 1.5 * torch.randn(B, D, device=dev)
    # parameter rows at the 16-byte-multiple pitch the conditioner's final layer writes them with (324 floats)
    cond = ops._padded_empty(B, P, torch.float32, dev)
    cond.copy_(0.5 * torch.randn(B, P, device=dev))
    gcond = ops._padded_empty(B, P, torch.float32, dev)
    shared = layer.unconditional_transform.packed_params().detach().contiguous()
    gy, gld = torch.randn(B, D, device=dev), torch.randn(B, device=dev)
    y, ld, gx = torch.empty_like(x), torch.empty(B, device=dev), torch.empty_like(x)
    gshared = torch.empty_like(shared)
    ws = torch.empty(_lib.query("vcnf_rqs_layer_bwd_workspace", spec.Ts, spec.Ps, 0), dtype=torch.uint8, device=dev)
    cmap, cm, cb, sm, sb = spec.on(dev, torch.float32)
    peak, peak_src = _peak()
    out = {}

    def fwd():              # straight through the C ABI (no autograd bookkeeping between the events)
        _lib.call("vcnf_rqs_layer_fwd", ptr(x), D, ptr(y), D, ptr(ld), 0, B, D, spec.K, ptr(cmap), ptr(cond),
                  cond.stride(0), spec.Tc, spec.Pc, ptr(cm), ptr(cb), spec.wh_scale, ptr(shared), spec.Ts, spec.Ps,
                  ptr(sm), ptr(sb), 0, spec.outside_zero, None, None, 0, torch.cuda.current_stream().cuda_stream)

    def bwd():              # layer backward + the fixed-order reduce of the batch-shared gradients (2 launches)
        _lib.call("vcnf_rqs_layer_bwd", ptr(x), D, ptr(gy), D, ptr(gld), ptr(gx), D, B, D, spec.K, ptr(cmap), ptr(cond),
                  cond.stride(0), spec.Tc, spec.Pc, ptr(cm), ptr(cb), spec.wh_scale, ptr(gcond), gcond.stride(0),
                  ptr(shared), spec.Ts, spec.Ps, ptr(sm), ptr(sb), ptr(gshared), ptr(ws), 0, spec.outside_zero, 0,
                  torch.cuda.current_stream().cuda_stream)

    P = spec.Tc * spec.Pc
    # conditioner dense layer at the hidden shape (64 of the 96 forward GEMMs of a step, and the same
    # kernel runs the 96 input-gradient GEMMs): y = relu(h) W^T + b + h
    h = torch.randn(B, HIDDEN, device=dev)
    W = torch.randn(HIDDEN, HIDDEN, device=dev) * 0.05
    bvec = torch.randn(HIDDEN, device=dev)
    gh = torch.randn(B, HIDDEN, device=dev)

    def lin_fwd():
        with torch.no_grad():
            ops.linear(h, W, bvec, residual=h, relu_in=True)

    gW, gbv = torch.empty_like(W), torch.empty_like(bvec)
    wws = torch.empty(_lib.query("vcnf_linear_wgrad_workspace", B, HIDDEN, HIDDEN), dtype=torch.uint8, device=dev)

    def lin_wgrad():        # straight through the C ABI (no autograd bookkeeping between the events)
        _lib.call("vcnf_linear_wgrad_tf32x3", gh.data_ptr(), HIDDEN, h.data_ptr(), HIDDEN, gW.data_ptr(),
                  gbv.data_ptr(), wws.data_ptr(), B, HIDDEN, HIDDEN, 1, torch.cuda.current_stream().cuda_stream)

    kernels = (("linear_ts_kernel (vcnf_linear_tf32x3 fwd/dgrad: hidden 128x128, relu+bias+residual fused)", lin_fwd,
                4 * (3 * HIDDEN), 224),                       # read h (A and residual), write y
               ("linear_wgrad_tma_kernel (vcnf_linear_wgrad_tf32x3, + reduce)", lin_wgrad, 4 * (2 * HIDDEN), 96),
               ("rqs_layer_fwd_fast_kernel (vcnf_rqs_layer_fwd)", fwd, 4 * (P + 2 * D + 1), 16),
               ("rqs_layer_bwd_fast_kernel (vcnf_rqs_layer_bwd, + shared-grad reduce)", bwd, 4 * (2 * P + 3 * D + 1), 16))
    for name, fn, bytes_per_sample, per_step in kernels:
        ms = _time_launches(fn, flush)
        ach = bytes_per_sample * B / (ms * 1e-3) / 1e9
        out[name] = {"ms": ms, "achieved": ach, "frac": ach / peak, "bytes_per_sample": bytes_per_sample,
                     "launches_per_step": per_step, "step_share_ms": ms * per_step}
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        measured = json.load(open(tpath)) if os.path.exists(tpath) else {}
    except Exception:
        measured = {}
    for name in out:        # dram bytes per launch from the committed `ncu --set full` captures
        out[name]["traffic"] = measured.get(name.split(" ")[0])
    dom_name = max(out, key=lambda k: out[k]["step_share_ms"])
    dom = out[dom_name]
    traffic = dom["traffic"]
    return {"bound": "hbm", "kernel": dom_name,
            "achieved": dom["achieved"], "peak": peak, "peak_source": peak_src, "unit": "GB/s",
            "frac": dom["frac"], "traffic": traffic, "algorithmic_bytes_per_launch": dom["bytes_per_sample"] * B,
            "launch_ms": dom["ms"],
            "note": "dominant = largest (launch time x launches per step); the fp32-faithful GEMM moves "
                    "activations once in / once out, so its roofline is HBM (SURVEY.md 8d)",
            "other_kernels": {k: v for k, v in out.items() if k != dom_name}}


def gpu_run(args):
    import vcnf_b200 as nf
    from vcnf_b200 import ops, parallel
    import torch.distributed as dist

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path (use --impl reference)")
    rank, world, local = parallel.init_from_env()
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    torch.manual_seed(0)
    model = build_model(nf).to(dev)
    if args.torch_adam:
        bucket = parallel.GradBucket(model.parameters())
        opt = torch.optim.Adam(model.parameters(), lr=1e-3, fused=True, capturable=not args.no_graph)
    else:
        # Adam on flat parameter / gradient / moment buffers, one launch per step (vcnf_b200/optim.py)
        opt = nf.optim.FlatAdam(model.parameters(), lr=1e-3)
        bucket = opt.bucket
    x_host = synth_batch(BATCH, 1234 + rank).pin_memory()
    x_dev = x_host.to(dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    if args.no_graph:
        def step(x):
            bucket.begin()
            loss = model.forward_kld(x)
            loss.backward()
            if world > 1:
                bucket.collect()
                bucket.allreduce_mean(world)
            opt.step()
            return loss
        graphed = False
    else:
        # the whole step (~1100 launches) captured once into a CUDA graph and replayed (parallel.py)
        step = parallel.GraphedTrainStep(model, opt, bucket, x_dev, world=world)
        graphed = step.graph is not None
    launches_per_step = None

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step(x_dev)
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    if not graphed:
        launches0 = ops.LAUNCHES
    else:
        # kernels inside one replay = kernels the eager step launches (counted on one eager run)
        launches0 = ops.LAUNCHES
        step._eager()
        launches_per_step = ops.LAUNCHES - launches0
        barrier()
    evs = []
    barrier()
    for _ in range(args.steps):
        flush.zero_()                                   # L2 flush, outside the per-step event pair
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        loss = step(x_dev)
        e1.record()
        evs.append((e0, e1))
    barrier()
    launches = (ops.LAUNCHES - launches0) if not graphed else launches_per_step * args.steps
    total_ms = sum(a.elapsed_time(b) for a, b in evs)
    clocks = sampler.summary()
    t = torch.tensor([total_ms], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t)
    value = world * BATCH * args.steps / (total_ms * 1e-3)

    # end to end through the public API: pinned host batch -> device every step, loss read back
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        xb = x_host.to(dev, non_blocking=True)
        loss_val = step(xb).item()
    barrier()
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = world * BATCH * args.steps / float(t)

    # sampling direction (second half of the BASELINE metric), device-timed
    draw = parallel.GraphedSampler(model, BATCH) if not args.no_graph else (lambda: model.sample(BATCH))
    with torch.no_grad():
        for _ in range(3):
            draw()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            flush.zero_()               # L2 flush between draws, as between training steps
            z_s, lq_s = draw()
        e1.record()
        torch.cuda.synchronize()
        ef0, ef1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ef0.record()
        for _ in range(10):
            flush.zero_()
        ef1.record()
        torch.cuda.synchronize()
        sample_ms = (e0.elapsed_time(e1) - ef0.elapsed_time(ef1)) / 10
        sample_rate = world * BATCH / (sample_ms * 1e-3)
        sample_ok = bool(torch.isfinite(z_s).all() and torch.isfinite(lq_s).all())

    if rank != 0:
        return
    roof = kernel_roofline(model, dev, flush)
    cpu = cpu_reference_run(steps=2, warmup=1) if world == 1 else None
    line = {
        "metric": "forward_kld_train_samples_per_sec", "value": value, "unit": "samples/s",
        "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "global_batch": world * BATCH, "features": D, "layers": LAYERS,
                   "hidden": HIDDEN, "bins": BINS, "parallelism": "dp%d" % world,
                   "l2": "256 MiB buffer written between timed steps (L2 flush)",
                   "conditioner_gemms": "tcgen05 3xTF32 (vcnf_linear_tf32x3 fwd/dgrad, vcnf_linear_wgrad_tf32x3)", "optimizer": "torch Adam(fused)" if args.torch_adam else "Adam, flat buffers (vcnf_adam_flat)", "cuda_graph": graphed,
                   "cuda_graph_error": getattr(step, "capture_error", None) if not args.no_graph else None},
        "clocks": clocks, "gpu_launches": launches,
        "e2e": {"value": e2e_value, "unit": "samples/s", "h2d_bytes_per_step": BATCH * D * 4,
                "d2h_bytes_per_step": 4},
        "roofline": roof, "cpu_baseline": cpu,
        "sample": {"value": sample_rate, "unit": "samples/s", "batch": BATCH, "ms_per_draw": sample_ms,
                   "cuda_graph": bool(getattr(draw, "graph", None) is not None),
                   "cuda_graph_error": getattr(draw, "capture_error", None), "finite": sample_ok},
        "loss": float(loss_val),
    }
    _emit(line)


def reference_run(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    steps, warmup = max(1, min(args.steps, 5)), max(1, min(args.warmup, 2))
    r = cpu_reference_run(steps=steps, warmup=warmup)
    world = int(os.environ.get("WORLD_SIZE", str(args.gpus)))
    line = {
        "impl": "reference", "metric": "forward_kld_train_samples_per_sec", "value": r["value"],
        "unit": "samples/s", "n_gpus": world, "steps": steps, "warmup": warmup,
        "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "global_batch": CPU_SAMPLE, "features": D, "layers": LAYERS,
                   "hidden": HIDDEN, "bins": BINS, "parallelism": "cpu"},
        "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
        "e2e": {"value": r["value"], "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    _emit(line)


_JSON_FD = None


def _guard_stdout():
    """Keep stdout for the ONE JSON line: everything else written to fd 1 by libraries (NCCL prints its
    version banner there) is routed to stderr."""
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)


def _emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def main():
    _guard_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-graph", action="store_true", help="run the step eagerly instead of replaying a CUDA graph")
    ap.add_argument("--torch-adam", action="store_true", help="torch.optim.Adam(fused=True) instead of optim.FlatAdam")
    args = ap.parse_args()
    if args.impl == "reference":
        reference_run(args)
    else:
        gpu_run(args)
    import torch.distributed as dist
    if dist.is_initialized():
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
