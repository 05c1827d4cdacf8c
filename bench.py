#!/usr/bin/env python
"""Benchmark of the bijector hot path (BASELINE.json metric: forward_kld train samples/s).

Workload (BASELINE.json configs[1], SURVEY.md 8d cfg 2): Neural Spline Flow,
16 x [CoupledRationalQuadraticSpline(29, 2 blocks, 128 hidden, 8 bins, linear tails, bound 3) +
LULinearPermute(29)], base DiagGaussian(29), batch 65 536 per GPU, fp32, Adam(1e-3); one "step"
= forward_kld + backward + (gradient all-reduce for N > 1) + optimizer step on one synthetic batch
x = 1.5 * randn(B, 29).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--config cfg2]

Prints ONE JSON line (rank 0).  `--impl reference` times the CPU oracle port of the reference
algorithm (oracle/, torch CPU, all host threads) on the same workload (bounded number of steps).

`--config {cfg1,cfg2,cfg3a,cfg3b,cfg4,cfg5}` selects another BASELINE.json configuration for the whole
line; the default line (cfg 2) additionally carries a `configs` object with a short measurement of each of the
others (train step and, where the model samples, sample()), so that one driver run records all of them.
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
def _cpu_rows():
    """Rows per CPU step: the full 65 536-row batch when the host has the memory for the oracle's autograd graph
    (~0.6 MB per row over 16 layers), else a quarter of it."""
    try:
        import psutil
        avail = psutil.virtual_memory().available
    except Exception:
        avail = 0
    return BATCH if avail > 64 * (1 << 30) else (BATCH // 4 if avail > 20 * (1 << 30) else CPU_SAMPLE)


def cpu_reference_run(steps, warmup, sample=CPU_SAMPLE):
    """The reference algorithm's CPU path: oracle port, all host threads.  Warm-up steps run on 1 024 rows (thread pool,
    allocator); the timed steps on `sample` rows of the bench batch."""
    import oracle
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    torch.manual_seed(0)
    model = build_model(oracle)
    opt = torch.optim.Adam(model.parameters(), lr=1e-3)
    x = synth_batch(BATCH, 1234)[:sample]
    times = []
    for i in range(warmup + steps):
        xb = x[:1024] if i < warmup else x
        t0 = time.perf_counter()
        opt.zero_grad(set_to_none=True)
        loss = model.forward_kld(xb.clone())
        loss.backward()
        opt.step()
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
    total = sum(times)
    return {"value": sample * len(times) / total, "unit": "samples/s", "cores": cores, "kind": "port",
            "sample": "%d of the %d-row batch per step, %d timed step(s) after %d warm-up step(s) on 1024 rows, "
                      "torch CPU %d threads" % (sample, BATCH, len(times), warmup, torch.get_num_threads()),
            "rows_per_step": sample, "ms_per_step": 1e3 * total / len(times), "loss": float(loss.detach())}


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
                4 * (3 * HIDDEN), 128),                       # read h (A and residual), write y; 128 per step = the
                                                              # input-gradient GEMMs (the forward ones are inside
                                                              # resnet_spline_kernel since round 2, see other_kernels)
               ("linear_wgrad_tma_kernel (vcnf_linear_wgrad_tf32x3, + reduce)", lin_wgrad, 4 * (2 * HIDDEN), 96),
               ("rqs_layer_fwd_fast_kernel (vcnf_rqs_layer_fwd)", fwd, 4 * (P + 2 * D + 1), 0),   # not in the cfg-2 step any more
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


# --------------------------------------------------------------------------------- the other BASELINE configurations
CFG4_CATS = (1, 3, 4, 5, 6, 7, 8, 9, 13, 14)
CFG4_LEVELS = (9, 16, 16, 7, 15, 6, 5, 2, 42, 2)


def config_spec(name, nf):
    """SURVEY.md 8d instantiation of BASELINE.json `configs`: dict(build, batch, data(n, seed), loss_fn | None,
    optimizer(params), dtype, workload, sample_batch | None, setup)."""
    from cases import build_cfg4_small, cfg4_data, build_glow_small
    if name == "cfg1":
        def build():
            flows = []
            for _ in range(16):
                flows += [nf.flows.AffineCouplingBlock(nf.nets.MLP([1, 32, 32, 2], init_zeros=True)),
                          nf.flows.Permute(2, mode="swap")]
            return nf.NormalizingFlow(nf.distributions.DiagGaussian(2), flows, nf.distributions.TwoMoons())
        return dict(build=build, batch=1024, data=lambda n, seed: torch.zeros(1), dtype="f32",
                    loss_fn=lambda m, x, B=1024: m.reverse_kld(num_samples=B), lr=1e-4, wd=0.0, sample_batch=65536,
                    workload="cfg1 README 2-D Real NVP 16x[AffineCouplingBlock(MLP[1,32,32,2])+Permute swap], TwoMoons, "
                             "reverse_kld batch 1024/GPU, fp32")
    if name in ("cfg3a", "cfg3b"):
        def build():
            flows = []
            for _ in range(8):
                if name == "cfg3a":
                    flows += [nf.flows.AutoregressiveRationalQuadraticSpline(64, 2, 256), nf.flows.LULinearPermute(64)]
                else:
                    flows += [nf.flows.Reverse(nf.flows.MaskedAffineAutoregressive(64, 256)),
                              nf.flows.Permute(64, mode="shuffle")]
            return nf.NormalizingFlow(nf.distributions.DiagGaussian(64, trainable=False), flows,
                                      categoricals=[], catlevels=[], catvdeqs=[])
        layer = "AutoregressiveRQS(64,2,256)+LULinearPermute(64)" if name == "cfg3a" else \
            "Reverse(MaskedAffineAutoregressive(64,256))+Permute(64,shuffle)"
        return dict(build=build, batch=32768, dtype="f32", loss_fn=None, lr=1e-3, wd=0.0, sample_batch=32768,
                    data=lambda n, seed: 1.5 * torch.randn(n, 64, generator=torch.Generator().manual_seed(seed)),
                    workload="%s 8x[%s] forward_kld train, batch 32768/GPU (262144 over 8), fp32" % (name, layer))
    if name == "cfg4":
        return dict(build=lambda: build_cfg4_small(nf, d=15, cats=CFG4_CATS, levels=CFG4_LEVELS, n_main=12,
                                                   n_vdeq=8).double(),
                    batch=2048, dtype="f64", loss_fn=None, lr=1e-4, wd=1e-6, sample_batch=None,
                    data=lambda n, seed: cfg4_data(n=n, d=15, cats=CFG4_CATS, levels=CFG4_LEVELS, seed=seed),
                    workload="cfg4 runadultvdeq1ft recipe: 10 categorical columns x VariationalDequantization(8x[ActNorm,"
                             "MaskedAffineFlow(MLP[1,4,1]),ActNorm]) + 12x[MaskedAffineFlow(MLP[15,30,15]x2),ActNorm(15)], "
                             "GaussianMixture(100,15), fp64 model / fp32 batch 2048/GPU, forward_kld train")
    if name == "cfg5":
        def setup():
            from vcnf_b200.nets import cnn
            cnn.set_precision("bf16")
        return dict(build=lambda: build_glow_small(nf, L=3, K=16, hidden=256, in_shape=(3, 32, 32), classes=10),
                    batch=128, dtype="bf16 conv conditioner (fp32 accumulation), fp32 bijectors", loss_fn="xy",
                    lr=1e-3, wd=1e-5, sample_batch=None, setup=setup, optimizer="adamax",
                    data=lambda n, seed: (torch.rand(n, 3, 32, 32, generator=torch.Generator().manual_seed(seed))
                                          * 255 / 256 + torch.rand(n, 3, 32, 32,
                                                                   generator=torch.Generator().manual_seed(seed + 1)) / 256),
                    workload="cfg5 Glow multiscale L=3 K=16 hidden 256 (example/glow.ipynb), 3x32x32, 10 classes, "
                             "forward_kld train, batch 128/GPU, ConvNet2d on tcgen05 bf16 implicit-GEMM kernels")
    raise ValueError(name)


def measure_config(name, dev, world, rank, steps, warmup, flush, batch=None, with_e2e=True):
    """Train-step (and sample()) rate of one configuration, the step replayed from a CUDA graph."""
    import vcnf_b200 as nf
    from vcnf_b200 import ops, parallel
    import torch.distributed as dist
    spec = config_spec(name, nf)
    if spec.get("setup"):
        spec["setup"]()
    B = batch or spec["batch"]
    torch.manual_seed(0)
    model = spec["build"]().to(dev)
    x_host = spec["data"](B, 1234 + rank)
    x_host = x_host.pin_memory() if x_host.numel() > 1 else x_host
    x_dev = x_host.to(dev)
    loss_fn = spec["loss_fn"]
    if name == "cfg1":
        loss_fn = lambda m, x, B=B: m.reverse_kld(num_samples=B)       # noqa: E731
    if loss_fn == "xy":
        y = torch.randint(0, 10, (B,), generator=torch.Generator().manual_seed(99 + rank)).to(dev)
        loss_fn = lambda m, xb, y=y: m.forward_kld(xb, y)              # noqa: E731
    if spec.get("optimizer") == "adamax":
        # data-dependent ActNorm initialisation first (one eager step's forward), then the capturable optimiser
        with torch.no_grad():
            loss_fn(model, x_dev.clone())
        opt = nf.optim.FlatAdamax(model.parameters(), lr=spec["lr"], weight_decay=spec["wd"])
        bucket = opt.bucket
        opt_name = "Adamax, flat buffers (vcnf_adamax_flat_guarded)"
    else:
        if name == "cfg4":
            with torch.no_grad():
                model.forward_kld(x_dev.clone())                       # ActNorm initialisation (first call)
        opt = nf.optim.FlatAdam(model.parameters(), lr=spec["lr"], weight_decay=spec["wd"])
        bucket = opt.bucket
        opt_name = "Adam, flat buffers (vcnf_adam_flat_guarded)"
    step = parallel.GraphedTrainStep(model, opt, bucket, x_dev, world=world, loss_fn=loss_fn)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
    for _ in range(max(warmup, 3)):
        step(x_dev)
    barrier()
    l0 = ops.LAUNCHES
    step._eager()
    per_step = ops.LAUNCHES - l0
    barrier()
    evs = []
    for _ in range(steps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        loss = step(x_dev)
        e1.record()
        evs.append((e0, e1))
    barrier()
    t = torch.tensor([sum(a.elapsed_time(b) for a, b in evs)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t)
    out = {"workload": spec["workload"] if batch is None else spec["workload"] + " [batch %d/GPU]" % B,
           "batch_per_gpu": B, "ms_per_step": total_ms / steps, "value": world * B * steps / (total_ms * 1e-3),
           "unit": "samples/s", "steps": steps, "gpu_launches_per_step": per_step, "optimizer": opt_name,
           "cuda_graph": step.graph is not None, "cuda_graph_error": step.capture_error, "dtype": spec["dtype"],
           "loss": float(loss)}
    if with_e2e:
        barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            xb = x_host.to(dev, non_blocking=True)
            step(xb).item()
        barrier()
        t = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        out["e2e"] = {"value": world * B * steps / float(t), "unit": "samples/s",
                      "h2d_bytes_per_step": int(x_host.numel() * x_host.element_size()) if x_host.numel() > 1 else 0,
                      "d2h_bytes_per_step": 4}
    if spec.get("sample_batch"):
        n = spec["sample_batch"]
        draw = parallel.GraphedSampler(model, n)
        for _ in range(2):
            draw()
        torch.cuda.synchronize()
        reps = 5
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            z_s, lq_s = draw()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / reps
        out["sample"] = {"value": world * n / (ms * 1e-3), "unit": "samples/s", "batch": n, "ms_per_draw": ms,
                         "cuda_graph": draw.graph is not None, "finite": bool(torch.isfinite(z_s).all())}
        del draw
    del step, opt, bucket, model
    torch.cuda.empty_cache()
    return out


def extra_kernel_rooflines(dev, flush):
    """HBM roofline of the affine family, the per-row linear maps and the optimiser kernel at their configuration
    shapes (SURVEY.md 8d algorithmic bytes), straight through the C ABI, CUDA-event timed after an L2 flush."""
    from vcnf_b200 import _lib, ops
    from vcnf_b200._lib import ptr
    peak, _ = _peak()
    st = lambda: torch.cuda.current_stream().cuda_stream       # noqa: E731
    out = {}

    def add(name, fn, bytes_per_launch, note=None):
        ms = _time_launches(fn, flush, reps=10)
        ach = bytes_per_launch / (ms * 1e-3) / 1e9
        out[name] = {"ms": ms, "achieved": ach, "frac": ach / peak, "unit": "GB/s", "bytes_per_launch": bytes_per_launch}
        if note:
            out[name]["note"] = note
    # K1 affine coupling block, cfg 1 rows at the roofline-meaningful batch 2^20: 28 B / sample fwd, 52 B bwd
    B = 1 << 20
    z, prm = torch.randn(B, 2, device=dev), 0.1 * torch.randn(B, 2, device=dev)
    o, ld, gz, gp = torch.empty_like(z), torch.empty(B, device=dev), torch.empty_like(z), torch.empty_like(prm)
    add("coupling_kernel<fwd> (vcnf_affine_coupling_fwd, cfg 1, B = 2^20)",
        lambda: _lib.call("vcnf_affine_coupling_fwd", ptr(z), ptr(prm), ptr(o), ptr(ld), B, 2, 1, 1, 1, 0, 0, 0, 0, st()),
        28 * B)
    add("coupling_kernel<bwd> (vcnf_affine_coupling_bwd, cfg 1, B = 2^20)",
        lambda: _lib.call("vcnf_affine_coupling_bwd", ptr(z), ptr(prm), ptr(o), ptr(ld), ptr(gz), ptr(gp), B, 2, 1, 1, 1,
                          0, 0, 0, 0, st()), 52 * B)
    # K2 masked affine, cfg 4 main flow (d = 15, fp64) at B = 65536: s (4 D + 1) fwd, s (7 D + 1) bwd
    B, n = 65536, 15
    z = torch.randn(B, n, device=dev, dtype=torch.float64)
    sc, tr = 0.1 * torch.randn_like(z), torch.randn_like(z)
    mask = (torch.arange(n, device=dev) % 2).double()
    o, ld = torch.empty_like(z), torch.empty(B, device=dev, dtype=torch.float64)
    gz, gs, gt = torch.empty_like(z), torch.empty_like(z), torch.empty_like(z)
    add("masked_affine_kernel<fwd> (vcnf_masked_affine_fwd, cfg 4, d = 15 fp64, B = 65536)",
        lambda: _lib.call("vcnf_masked_affine_fwd", ptr(z), ptr(mask), ptr(sc), ptr(tr), ptr(o), ptr(ld), B, n, 1, 1, st()),
        8 * (4 * n + 1) * B)
    add("masked_affine_kernel<bwd> (vcnf_masked_affine_bwd)",
        lambda: _lib.call("vcnf_masked_affine_bwd", ptr(z), ptr(mask), ptr(sc), ptr(tr), ptr(o), ptr(ld), ptr(gz), ptr(gs),
                          ptr(gt), B, n, 1, 1, st()), 8 * (7 * n + 1) * B)
    # K3 ActNorm (d = 15 fp64): s 2 D fwd, s 3 D bwd (+ parameter-gradient reduce)
    s_, t_ = torch.randn(n, device=dev, dtype=torch.float64) * 0.1, torch.randn(n, device=dev, dtype=torch.float64)
    gs_, gt_ = torch.empty_like(s_), torch.empty_like(t_)
    ws = torch.empty(_lib.query("vcnf_const_affine_bwd_workspace", n, 1), dtype=torch.uint8, device=dev)
    add("const_affine_kernel (vcnf_const_affine_fwd, ActNorm(15) fp64, B = 65536)",
        lambda: _lib.call("vcnf_const_affine_fwd", ptr(z), ptr(s_), ptr(t_), ptr(o), B, n, 1, 1, 1, st()), 8 * 2 * n * B)
    add("const_affine_bwd_kernel + reduce (vcnf_const_affine_bwd)",
        lambda: _lib.call("vcnf_const_affine_bwd", ptr(z), ptr(s_), ptr(t_), ptr(o), ptr(gz), ptr(gs_), ptr(gt_), ptr(ws),
                          B, n, 1, 1, 1, st()), 8 * 3 * n * B)
    # K10 LULinearPermute rows (cfg 2: D = 29 fp32, B = 65536): s (2 D + 1) fwd, s 3 D bwd
    B, n = BATCH, D
    x, W = torch.randn(B, n, device=dev), torch.randn(n, n, device=dev) / 5
    bias, y, gx = torch.randn(n, device=dev), torch.empty(B, n, device=dev), torch.empty(B, n, device=dev)
    perm = torch.randperm(n, device=dev).to(torch.int32)
    gW, gb = torch.empty_like(W), torch.empty_like(bias)
    ws2 = torch.empty(_lib.query("vcnf_small_linear_bwd_workspace", n, 0), dtype=torch.uint8, device=dev)
    add("small_linear_rows_kernel (vcnf_small_linear, LULinearPermute(29), B = 65536)",
        lambda: _lib.call("vcnf_small_linear", ptr(x), ptr(y), ptr(W), ptr(bias), ptr(perm), None, B, n, 1, 0, 0, st()),
        4 * (2 * n + 1) * B)
    add("small_linear_bwd_kernel + reduce (vcnf_small_linear_bwd)",
        lambda: _lib.call("vcnf_small_linear_bwd", ptr(x), ptr(y), ptr(W), ptr(perm), None, ptr(gx), ptr(gW), ptr(gb),
                          ptr(ws2), B, n, 0, st()), 4 * 3 * n * B)
    lo, up, dg = torch.randn(n * (n - 1) // 2, device=dev) / 5, torch.randn(n * (n - 1) // 2, device=dev) / 5, torch.randn(n, device=dev)
    lad = torch.empty((), device=dev)
    add("lu_compose_fwd_kernel (vcnf_lu_compose_fwd, D = 29)",
        lambda: _lib.call("vcnf_lu_compose_fwd", ptr(lo), ptr(up), ptr(dg), 1e-3, n, ptr(W), ptr(lad), 0, st()),
        4 * (n * n + n * n), note="launch-bound single-CTA kernel: D^2 parameters; the time is what matters")
    # optimiser: 28 B per fp32 parameter (cfg 2: 1.8 M parameters)
    N = 1835008
    p_, g_, m_, v_ = (torch.randn(N, device=dev) for _ in range(4))
    v_.abs_()
    stp, tick = torch.zeros(1, device=dev), torch.zeros(1, dtype=torch.int32, device=dev)
    add("adam_flat_kernel (vcnf_adam_flat, 1.8 M fp32 parameters)",
        lambda: _lib.call("vcnf_adam_flat", ptr(p_), ptr(g_), ptr(m_), ptr(v_), ptr(stp), ptr(tick), N, 1e-3, 0.9, 0.999,
                          1e-8, 0.0, 0, st()), 28 * N)
    # K6 register-resident MLP (cfg 1 [1,32,32,2], B = 2^20): FP32 FMA pipe, 2 240 flop / sample
    B = 1 << 20
    xm = torch.randn(B, 1, device=dev)
    mlp = [torch.randn(32, 1, device=dev), torch.randn(32, 32, device=dev) / 6, torch.randn(2, 32, device=dev) / 6]
    mb = [torch.randn(32, device=dev), torch.randn(32, device=dev), torch.randn(2, device=dev)]
    with torch.no_grad():
        ms = _time_launches(lambda: ops.mlp_small(xm, mlp, mb, 0.0), flush, reps=10)
    flops = 2.0 * (32 + 32 * 32 + 64) * B
    out["mlp_fwd_kernel (vcnf_mlp_fwd, cfg 1 [1,32,32,2], B = 2^20)"] = {
        "ms": ms, "achieved": 12 * B / (ms * 1e-3) / 1e9, "frac": 12 * B / (ms * 1e-3) / 1e9 / peak, "unit": "GB/s",
        "bytes_per_launch": 12 * B, "tflops_fp32_fma": flops / (ms * 1e-3) / 1e12,
        "note": "FP32 FMA-pipe bound (~100 flop / B, SURVEY.md K6): the TFLOP/s figure is the relevant one"}
    # fused conditioner + spline (vcnf_resnet_spline_fwd), one cfg-2 layer in the no-grad direction on the bench batch:
    # tensor-bound -- 3 TF32 MMAs per product; peak = half the measured bf16 rate (TF32 runs at half the bf16 rate)
    try:
        import vcnf_b200 as nf
        layer = nf.flows.CoupledRationalQuadraticSpline(D, BLOCKS, HIDDEN, num_bins=BINS).to(dev)
        c = layer.prqct
        xb = torch.randn(BATCH, D, device=dev)
        with torch.no_grad():
            fused = c._fused_route(xb, None)
            w, bs = c.transform_net.fused_params()
            shared = c._shared_params()
            packed = fused.pack(w, bs)
            fused.pack = lambda *_: packed                     # time the kernel, not the 4 us re-tiling launch
            ms = _time_launches(lambda: ops.resnet_spline(xb, fused, w, bs, inverse=False, shared=shared), flush, reps=10)
        n_id = len(c.identity_features)
        macs = n_id * HIDDEN + 2 * BLOCKS * HIDDEN * HIDDEN + HIDDEN * fused.Tc * fused.Pc      # per row, unpadded
        try:
            tpeak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops"]) / 2
            tsrc = "MEASURED_PEAKS.json bf16_tflops / 2 (TF32 = half the bf16 rate)"
        except Exception:
            tpeak, tsrc = 1100.0, "fallback (B200_PROFILING.md: 1.1 PFLOP/s dense TF32)"
        tf = 3 * 2.0 * macs * BATCH / (ms * 1e-3) / 1e12
        out["resnet_spline_kernel (vcnf_resnet_spline_fwd, cfg 2 layer, no-grad direction)"] = {
            "ms": ms, "bound": "tensor", "achieved": tf, "unit": "TFLOP/s", "peak": tpeak, "peak_source": tsrc,
            "frac": tf / tpeak, "flops_per_launch": 3 * 2.0 * macs * BATCH,
            "bytes_per_launch": (2 * 4 * D + 4) * BATCH,
            "note": "executed TF32 flops (3 MMAs per fp32 product, unpadded shapes); HBM traffic is x in, y and log|det| out"}
        try:        # dram bytes of one launch from the committed ncu --set full capture (profiles/r2_fused_ncu_summary.txt:
            #         the 7.6 MB of y were still in L2 when the capture ended, so this is the read side)
            out["resnet_spline_kernel (vcnf_resnet_spline_fwd, cfg 2 layer, no-grad direction)"]["traffic"] = json.load(
                open(os.path.join(ROOT, "profiles", "traffic.json"))).get("resnet_spline_kernel")
        except Exception:                                      # noqa: BLE001
            pass
    except Exception as exc:                                   # noqa: BLE001
        out["resnet_spline_kernel"] = {"error": repr(exc)[:200]}
    return out


def conv_roofline(dev, flush):
    """Tensor-pipe fraction of the cfg-5 convolution kernels at the level-2 shapes (B = 128, 16x16, hidden 256), live."""
    from vcnf_b200 import ops
    B, H, W, Hd, C = 128, 16, 16, 256, 12
    bf = torch.bfloat16
    h = torch.randn(B * H * W, Hd, device=dev).to(bf)
    w1 = (torch.randn(1, Hd, Hd, device=dev) / 16).to(bf)
    w9 = (torch.randn(9, 16, Hd, device=dev) / 48).to(bf)
    bias = torch.zeros(Hd, device=dev)
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        peak = float(json.load(open(path))["bf16_tflops_sustained"])
        src = "measured (MEASURED_PEAKS.json bf16_tflops_sustained)"
    except Exception:
        peak, src = 1400.0, "fallback (B200_PROFILING.md)"
    out = {}
    for label, fn, flops in (
            ("conv_gemm_bf16_kernel 1x1 256->256 @16x16 B=128", lambda: ops.conv_gemm_bf16(h, B, H, W, 1, w1, bias, epi=1),
             2.0 * B * H * W * Hd * Hd),
            ("conv_gemm_bf16_kernel 3x3 256->12 @16x16 B=128", lambda: ops.conv_gemm_bf16(h, B, H, W, 9, w9, None,
                                                                                          out_nchw_channels=C),
             2.0 * B * H * W * 9 * Hd * C),
            ("conv_wgrad_bf16_kernel 256x256 over 32768 pixels (+ reduce)", lambda: ops.conv_wgrad_bf16(h, h),
             2.0 * B * H * W * Hd * Hd)):
        ms = _time_launches(fn, flush, reps=10)
        tf = flops / (ms * 1e-3) / 1e12
        out[label] = {"ms": ms, "achieved": tf, "unit": "TFLOP/s", "frac": tf / peak, "flops_per_launch": flops}
    dom = max(out, key=lambda k: out[k]["ms"])
    return {"bound": "tensor", "kernel": dom, "achieved": out[dom]["achieved"], "peak": peak, "peak_source": src,
            "unit": "TFLOP/s", "frac": out[dom]["frac"], "traffic": None, "launch_ms": out[dom]["ms"],
            "note": "real (unpadded) multiply-adds of the convolution / kernel time; the 3x3 -> C layer pads C to 16 "
                    "columns and re-reads the activation tile once per tap from L2",
            "other_kernels": {k: v for k, v in out.items() if k != dom}}


def reference_eager_cuda(dev, steps=3):
    """The reference algorithm as users run it today: the oracle port (plain PyTorch ops, no package kernel) moved
    to the same B200, same model / batch / optimiser, eager, CUDA-event timed.  A baseline beside `cpu_baseline`."""
    try:
        import oracle
        torch.manual_seed(0)
        model = build_model(oracle).to(dev)
        opt = torch.optim.Adam(model.parameters(), lr=1e-3)
        x = synth_batch(BATCH, 1234).to(dev)

        def step():
            opt.zero_grad(set_to_none=True)
            loss = model.forward_kld(x.clone())
            loss.backward()
            opt.step()
            return loss
        step()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            loss = step()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        out = {"value": BATCH / (ms * 1e-3), "unit": "samples/s", "ms_per_step": ms, "steps": steps,
               "kind": "oracle port (.to('cuda')), eager PyTorch ops on the same GPU, fp32 (TF32 off)",
               "loss": float(loss)}
        del model, opt, x
        torch.cuda.empty_cache()
        return out
    except Exception as exc:            # noqa: BLE001
        return {"unavailable": repr(exc)[:300]}


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
    # (input pipeline of a training loop: while step i computes, the pinned batch of step i + 1 crosses PCIe on a copy
    # stream into the other of two device buffers; every step's copy and its loss read-back are inside the timed region)
    copy_stream = torch.cuda.Stream()
    stage = [torch.empty_like(x_dev), torch.empty_like(x_dev)]
    ready = [torch.cuda.Event(), torch.cuda.Event()]

    def prefetch(i):
        with torch.cuda.stream(copy_stream):
            stage[i & 1].copy_(x_host, non_blocking=True)
            ready[i & 1].record(copy_stream)

    barrier()
    t0 = time.perf_counter()
    prefetch(0)
    for i in range(args.steps):
        torch.cuda.current_stream().wait_event(ready[i & 1])
        loss_t = step(stage[i & 1])
        if i + 1 < args.steps:
            prefetch(i + 1)             # buffer (i + 1) & 1 was last read by step i - 1, whose loss has been read back
        loss_val = loss_t.item()
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
        # model.log_prob(x) on the bench batch, no gradients (evaluation): eager calls, CUDA events
        logp = parallel.GraphedLogProb(model, x_dev) if not args.no_graph else (lambda xb: model.log_prob(xb))
        for _ in range(3):
            logp(x_dev)
        torch.cuda.synchronize()
        l0, l1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0.record()
        for _ in range(10):
            lp = logp(x_dev)
        l1.record()
        torch.cuda.synchronize()
        logp_ms = l0.elapsed_time(l1) / 10
        logp_ok = bool(torch.isfinite(lp).all())
        logp_graph = bool(getattr(logp, "graph", None) is not None)

    # strong-scaling point of cfg 3a (BASELINE: "batch 256K sharded over 8 GPUs"): global batch fixed at 262 144
    strong = None
    step_err = getattr(step, "capture_error", None) if not args.no_graph else None
    draw_graph, draw_err = bool(getattr(draw, "graph", None) is not None), getattr(draw, "capture_error", None)
    if not args.no_configs:
        del step, draw
        torch.cuda.empty_cache()
        try:
            strong = measure_config("cfg3a", dev, world, rank, steps=5, warmup=3, flush=flush, batch=262144 // world,
                                    with_e2e=False)
            strong["scaling"] = "strong"
            strong["global_batch"] = 262144
            strong.pop("sample", None)
        except Exception as exc:        # noqa: BLE001
            strong = {"unavailable": repr(exc)[:300]}
    if rank != 0:
        return
    roof = kernel_roofline(model, dev, flush)
    others = {}
    eager_cuda = None
    cpu = None
    if world == 1:
        try:
            roof["other_kernels"].update(extra_kernel_rooflines(dev, flush))
        except Exception as exc:        # noqa: BLE001
            roof["other_kernels"]["error"] = repr(exc)[:300]
        if not args.no_configs:
            for name in ("cfg1", "cfg3a", "cfg3b", "cfg4", "cfg5"):
                try:
                    others[name] = measure_config(name, dev, 1, 0, steps=5, warmup=3, flush=flush)
                    if name == "cfg1":
                        big = measure_config(name, dev, 1, 0, steps=5, warmup=3, flush=flush, batch=1 << 20,
                                             with_e2e=False)
                        others[name]["batch_2^20"] = {k: big[k] for k in ("ms_per_step", "value", "unit")}
                    if name == "cfg4":
                        big = measure_config(name, dev, 1, 0, steps=5, warmup=3, flush=flush, batch=65536, with_e2e=False)
                        others[name]["batch_65536"] = {k: big[k] for k in ("ms_per_step", "value", "unit")}
                    if name == "cfg5":
                        big = measure_config(name, dev, 1, 0, steps=5, warmup=3, flush=flush, batch=512, with_e2e=False)
                        others[name]["batch_512"] = {k: big[k] for k in ("ms_per_step", "value", "unit")}
                        others[name]["roofline"] = conv_roofline(dev, flush)
                except Exception as exc:        # noqa: BLE001
                    others[name] = {"unavailable": repr(exc)[:300]}
            from vcnf_b200.nets import cnn
            cnn.set_precision("fp32")
        eager_cuda = reference_eager_cuda(dev)
        cpu = cpu_reference_run(steps=1, warmup=1, sample=_cpu_rows())
    line = {
        "metric": "forward_kld_train_samples_per_sec", "value": value, "unit": "samples/s",
        "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "global_batch": world * BATCH, "features": D, "layers": LAYERS,
                   "hidden": HIDDEN, "bins": BINS, "parallelism": "dp%d" % world,
                   "l2": "256 MiB buffer written between timed steps (L2 flush)",
                   "conditioner_gemms": "tcgen05 3xTF32 (vcnf_linear_tf32x3 fwd/dgrad, vcnf_linear_wgrad_tf32x3)", "optimizer": "torch Adam(fused)" if args.torch_adam else "Adam, flat buffers (vcnf_adam_flat)", "cuda_graph": graphed,
                   "cuda_graph_error": step_err},
        "clocks": clocks, "gpu_launches": launches,
        "e2e": {"value": e2e_value, "unit": "samples/s", "h2d_bytes_per_step": BATCH * D * 4,
                "d2h_bytes_per_step": 4,
                "note": "pinned host batch -> device every step on a copy stream (double-buffered: the copy of step i+1 "
                        "overlaps the compute of step i), loss.item() every step; wall clock over the whole loop"},
        "roofline": roof, "cpu_baseline": cpu, "reference_eager_cuda": eager_cuda,
        "configs": others, "cfg3a_strong": strong,
        "sample": {"value": sample_rate, "unit": "samples/s", "batch": BATCH, "ms_per_draw": sample_ms,
                   "cuda_graph": draw_graph, "cuda_graph_error": draw_err, "finite": sample_ok},
        "log_prob": {"value": world * BATCH / (logp_ms * 1e-3), "unit": "samples/s", "batch": BATCH, "ms_per_call": logp_ms,
                     "finite": logp_ok, "cuda_graph": logp_graph,
                     "note": "model.log_prob(x) without gradients on the bench batch (parallel.GraphedLogProb)"},
        "loss": float(loss_val),
    }
    _emit(line)


def config_run(args):
    """`--config X`: the whole line for another BASELINE configuration."""
    from vcnf_b200 import parallel
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path (use --impl reference)")
    rank, world, local = parallel.init_from_env()
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    sampler = ClockSampler(local)
    sampler.start()
    r = measure_config(args.config, dev, world, rank, steps=args.steps, warmup=max(args.warmup, 3), flush=flush)
    clocks = sampler.summary()
    if rank != 0:
        return
    if args.config == "cfg5":
        roof = conv_roofline(dev, flush)
    else:
        extras = extra_kernel_rooflines(dev, flush)
        key = {"cfg1": "coupling_kernel<fwd>", "cfg4": "masked_affine_kernel<fwd>"}.get(args.config)
        if key:
            name = next(k for k in extras if k.startswith(key))
            peak, src = _peak()
            d = extras.pop(name)
            roof = {"bound": "hbm", "kernel": name, "achieved": d["achieved"], "peak": peak, "peak_source": src,
                    "unit": "GB/s", "frac": d["frac"], "traffic": None, "launch_ms": d["ms"], "other_kernels": extras}
        else:
            import vcnf_b200 as nf
            torch.manual_seed(0)
            roof = kernel_roofline(build_model(nf).to(dev), dev, flush)
            roof["note"] = "cfg 3 shares the conditioner GEMM / spline kernels of cfg 2; measured at the cfg-2 shapes"
    line = {
        "metric": "forward_kld_train_samples_per_sec" if args.config != "cfg1" else "reverse_kld_train_samples_per_sec",
        "value": r["value"], "unit": "samples/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": r["dtype"], "data": "synthetic",
        "config": {"workload": r["workload"], "global_batch": world * r["batch_per_gpu"], "parallelism": "dp%d" % world,
                   "l2": "256 MiB buffer written between timed steps (L2 flush)", "optimizer": r["optimizer"],
                   "cuda_graph": r["cuda_graph"], "cuda_graph_error": r["cuda_graph_error"]},
        "clocks": clocks, "gpu_launches": r["gpu_launches_per_step"] * args.steps, "e2e": r.get("e2e"),
        "roofline": roof, "cpu_baseline": None, "sample": r.get("sample"), "loss": r["loss"],
    }
    _emit(line)


def reference_run(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    rows = _cpu_rows()
    steps, warmup = max(1, min(args.steps, 2 if rows == BATCH else 5)), max(1, min(args.warmup, 1))
    r = cpu_reference_run(steps=steps, warmup=warmup, sample=rows)
    world = int(os.environ.get("WORLD_SIZE", str(args.gpus)))
    line = {
        "impl": "reference", "metric": "forward_kld_train_samples_per_sec", "value": r["value"],
        "unit": "samples/s", "n_gpus": world, "steps": steps, "warmup": warmup,
        "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "global_batch": rows, "features": D, "layers": LAYERS,
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
    ap.add_argument("--config", default="cfg2", choices=["cfg1", "cfg2", "cfg3a", "cfg3b", "cfg4", "cfg5"])
    ap.add_argument("--no-configs", action="store_true", help="default line without the `configs` / `cfg3a_strong` extras")
    args = ap.parse_args()
    if args.impl == "reference":
        reference_run(args)
    elif args.config != "cfg2":
        config_run(args)
    else:
        gpu_run(args)
    import torch.distributed as dist
    if dist.is_initialized():
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
