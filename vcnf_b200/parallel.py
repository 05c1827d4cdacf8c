"""Data-parallel plumbing for `forward_kld` training: one process per GPU (torchrun), batch
sharded across ranks, ONE flat-bucket gradient all-reduce per step (NCCL over NVLink 5 /
NVSwitch on the GPU box, gloo in the CPU tests).  The reference has no distributed code at all
(SURVEY.md 2.2); samples are independent in every bijector, so this is the only collective the
path needs.  Sampling shards with no communication."""
import os

import torch
import torch.distributed as dist


def init_from_env(backend=None):
    """Initialise torch.distributed from RANK / WORLD_SIZE / MASTER_* (torchrun contract).
    Returns (rank, world_size, local_rank); single process when WORLD_SIZE is unset."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        kwargs = {}
        if backend == "nccl":
            torch.cuda.set_device(local)
            kwargs["device_id"] = torch.device("cuda", local)
        dist.init_process_group(backend=backend, rank=rank, world_size=world, **kwargs)
    return rank, world, local


def shard_bounds(n, rank, world):
    """Contiguous shard [lo, hi) of n samples for `rank` (first n % world ranks get one extra)."""
    base, extra = divmod(n, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


class GradBucket:
    """Flat gradient bucket: grads are views into one contiguous buffer, so the all-reduce is a
    single collective on 7-32 MB (latency-bound on NVSwitch)."""

    def __init__(self, params, align=1):
        """`align`: every parameter's slice starts at a multiple of `align` elements (pad elements stay zero);
        optim.FlatAdam lays its parameter / moment buffers out the same way."""
        self.params = [p for p in params if p.requires_grad]
        self.offsets, total = [], 0
        for p in self.params:
            self.offsets.append(total)
            total += (p.numel() + align - 1) // align * align
        self.total = total
        if not self.params:
            self.flat = None
            return
        p0 = self.params[0]
        self.flat = torch.zeros(total, dtype=p0.dtype, device=p0.device)
        for p, off in zip(self.params, self.offsets):
            if p.dtype != p0.dtype:
                raise TypeError("GradBucket needs a single parameter dtype")
            p.grad = self.flat[off:off + p.numel()].view_as(p)

    def zero(self):
        if self.flat is not None:
            self.flat.zero_()

    # Accumulating into pre-bound .grad views costs one tiny `add_` launch per parameter and backward
    # (AccumulateGrad adds into a defined .grad; 256 launches per cfg-2 step).  The cheaper protocol:
    # begin() drops the .grad references so autograd hands over the freshly computed gradient tensors,
    # collect() packs them into the flat buffer with ONE multi-tensor copy and re-binds the views
    # (only needed when a collective follows; a single process can step on the fresh tensors directly).
    def begin(self):
        for p in self.params:
            p.grad = None

    def collect(self):
        if self.flat is None:
            return
        views, grads, missing = [], [], []
        for p, off in zip(self.params, self.offsets):
            v = self.flat[off:off + p.numel()].view_as(p)
            if p.grad is None:
                missing.append(v)
            elif p.grad.data_ptr() != v.data_ptr():
                views.append(v)
                grads.append(p.grad)
            p.grad = v
        if views:
            torch._foreach_copy_(views, grads)
        if missing:
            torch._foreach_zero_(missing)

    def allreduce_mean(self, world=None, scale=True):
        """Sum the bucket over the ranks; scale=False leaves the division by `world` to the consumer
        (optim.FlatAdam folds it into its kernel: grad_scale = 1 / world)."""
        if self.flat is None or not dist.is_initialized():
            return
        world = dist.get_world_size() if world is None else world
        if world > 1:
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM)
            if scale:
                self.flat.div_(world)


def sync_model(model, src=0):
    """Make every rank start from rank `src`'s parameters AND buffers (initial weights, the random permutations of
    Permute / LULinearPermute, feature masks, ActNorm flags): ranks seeded differently would otherwise train
    different models whose gradients are averaged.  In-place broadcasts, so views (optim.FlatAdam's flat buffer) stay
    valid.  No-op for a single process."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return
    want_cuda = dist.get_backend() == "nccl"
    for t in list(model.parameters()) + list(model.buffers()):
        if t.numel() and t.is_cuda == want_cuda:       # host-side flags (ActNorm's init marker) are set per process
            dist.broadcast(t.data, src)


def check_equal_shards(n_local):
    """Data-parallel steps average the per-rank mean losses with equal weight; that is the global-batch gradient
    only when every rank holds the same number of rows.  Raises if the shard sizes differ."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return
    dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
    sizes = torch.tensor([n_local, -n_local], dtype=torch.int64, device=dev)
    dist.all_reduce(sizes, op=dist.ReduceOp.MAX)
    if int(sizes[0]) != -int(sizes[1]):
        raise ValueError("data-parallel forward_kld needs equal shards on every rank (got between %d and %d rows): "
                         "drop or pad the remainder" % (-int(sizes[1]), int(sizes[0])))


def sync_actnorm_init(model, src=0):
    """After the first (initialising) step make every rank use rank `src`'s data-dependent
    ActNorm statistics (flows/normalization.py:22-27 is per-process in the reference)."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return
    for m in model.modules():
        if hasattr(m, "data_dep_init_done") and hasattr(m, "s") and hasattr(m, "t"):
            dist.broadcast(m.s.data, src)
            dist.broadcast(m.t.data, src)


def _after_failed_capture(exc):
    """A capture that raised leaves the CUDA generator in capture mode (the next eager `torch.randn` raises
    "Offset increment outside graph capture"): a trivial capture that succeeds runs the generator's prologue and
    epilogue again and clears it.  Returns the error text, with a hint for the one failure a caller can fix."""
    text = repr(exc)
    torch.cuda.synchronize()
    try:
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            torch.zeros(1, device="cuda").add_(1)
        del g
        torch.cuda.synchronize()
    except Exception:                   # noqa: BLE001 -- best effort, the caller falls back to eager steps
        pass
    if "legacy stream" in text or "StreamCaptureImplicit" in text or "StreamCaptureInvalidated" in text:
        text += (" [hint: a loss / output of an earlier eager backward on the default stream is still alive; "
                 "it keeps the parameters' AccumulateGrad nodes, which are bound to the legacy stream, in use. "
                 "Drop those references (del loss) before building the graphed step.]")
    return text


class GraphedTrainStep:
    """One `forward_kld` training step (zero grads -> loss -> backward -> [gradient all-reduce] -> optimizer
    step) captured ONCE into a CUDA graph and replayed.  A cfg-2 step is ~1600 kernel launches from a
    Python loop (core.py:55-61 keeps the per-layer loop on the host); measured on B200 the host needs
    ~29 ms to issue them while the kernels take less, so the launch stream -- not the GPU -- bounds the
    step.  Replaying the captured graph removes the host from the loop; the arithmetic is identical.

        step = GraphedTrainStep(model, opt, bucket, example_batch)      # warms up 3 eager steps, captures
        loss = step(batch)          # copies `batch` into the static input, replays, returns the 0-dim loss

    The optimizer must be graph-capturable (e.g. torch.optim.Adam(..., capturable=True)).  If capture is
    not possible (a layer with a host-side branch such as ActNorm's first-call init), `.graph` stays None,
    `.capture_error` says why and the step runs eagerly -- same results, launch-bound.  `loss_fn(model, x)`
    replaces the loss, e.g. `lambda m, x: m.reverse_kld(num_samples=1024)` for the README's 2-D Real NVP
    (cfg 1; the base distribution's draw is captured and every replay advances the Philox offset).
    Tensors of an earlier eager backward on the default stream (a kept `loss`) must be released first: they
    hold the parameters' AccumulateGrad nodes, whose stream is the legacy stream, and capture cannot wait on it."""

    def __init__(self, model, optimizer, bucket, example, world=None, warmup=3, loss_fn=None):
        self.model, self.opt, self.bucket = model, optimizer, bucket
        self.world = world
        self.loss_fn = loss_fn if loss_fn is not None else (lambda m, x: m.forward_kld(x))
        self.x = example.clone()
        self.graph, self.loss, self.capture_error = None, None, None
        w = self._world()
        if w > 1:
            sync_model(model)                 # same weights, permutations and masks on every rank
            check_equal_shards(len(example))
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        # The warm-up iterations must not train: parameters, moments and the step counter are what the caller built
        # when the first real step runs.  FlatAdam skips its update (`frozen`); any other optimiser runs with lr = 0
        # and gets its state back afterwards.  (Data-dependent ActNorm initialisation still happens here, on the
        # example batch, exactly as it would in the first real step.)
        flat = hasattr(optimizer, "frozen")
        saved = None
        if flat:
            optimizer.frozen = True
        else:
            import copy
            saved = (copy.deepcopy(optimizer.state_dict()), [g["lr"] for g in optimizer.param_groups])
            for g in optimizer.param_groups:
                g["lr"] = 0.0 * g["lr"] if not torch.is_tensor(g["lr"]) else torch.zeros_like(g["lr"])
        # forward_kld dequantises categorical columns INTO its argument (core.py:54): every run starts from a fresh
        # copy of the batch, and the model sees a detached alias so that no autograd history sticks to the static input
        with torch.cuda.stream(side):
            for _ in range(warmup):
                self.x.copy_(example)
                self._eager()
        torch.cuda.current_stream().wait_stream(side)
        if flat:
            optimizer.frozen = False
        else:
            optimizer.load_state_dict(saved[0])
            for g, lr in zip(optimizer.param_groups, saved[1]):
                g["lr"] = lr
        self.x.copy_(example)
        torch.cuda.synchronize()
        try:
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                self.loss = self._eager()
            self.graph = g
        except Exception as exc:            # noqa: BLE001 -- report and run eagerly
            self.capture_error = _after_failed_capture(exc)
            self.graph = None

    def _world(self):
        return self.world if self.world is not None else (dist.get_world_size() if dist.is_initialized() else 1)

    def _eager(self):
        self.bucket.begin()
        loss = self.loss_fn(self.model, self.x.detach())
        loss.backward()
        world = self._world()
        flat = hasattr(self.opt, "grad_scale")
        if world > 1:
            self.bucket.collect()
            self.bucket.allreduce_mean(world, scale=not flat)     # FlatAdam divides by `world` inside its kernel
        if flat:
            # a NaN / Inf loss must not reach the parameters or the moments (runcred.py:249-250 tests it on the
            # host; a replayed graph cannot): the optimiser kernel reads the loss and skips the update
            self.opt.guard = loss.detach()
            self.opt.grad_scale = 1.0 / world
        self.opt.step()
        return loss

    def __call__(self, x):
        if hasattr(self.opt, "sync_hyper"):
            self.opt.sync_hyper()             # a changed learning rate reaches the replayed kernel
        self.x.copy_(x, non_blocking=True)
        if self.graph is None:
            return self._eager()
        self.graph.replay()
        return self.loss


class GraphedSampler:
    """`model.sample(num_samples)` (core.py:144-168: base draw + every flow in the sampling direction) captured
    ONCE into a CUDA graph and replayed.  Eagerly the call is ~500 kernel launches from the Python layer loop and
    is bound by the launch stream, not by the GPU (6 M samples/s at 65 536 on B200 whatever the kernels do);
    replayed, the host is out of the loop.  Every replay draws fresh base samples (the CUDA generator's Philox
    offset is advanced by graph replays) and sees the current parameters.

        draw = GraphedSampler(model, 65536)
        z, log_q = draw()        # static output tensors, overwritten by the next call -- clone to keep them

    If capture is not possible `.graph` stays None and the call runs eagerly (same results)."""

    def __init__(self, model, num_samples, warmup=3):
        self.model, self.n = model, int(num_samples)
        self.graph, self.out, self.capture_error = None, None, None
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side), torch.no_grad():
            for _ in range(warmup):
                model.sample(self.n)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        try:
            g = torch.cuda.CUDAGraph()
            with torch.no_grad(), torch.cuda.graph(g):
                self.out = model.sample(self.n)
            self.graph = g
        except Exception as exc:            # noqa: BLE001 -- report and run eagerly
            self.capture_error = _after_failed_capture(exc)
            self.graph = None

    def __call__(self):
        if self.graph is None:
            with torch.no_grad():
                return self.model.sample(self.n)
        self.graph.replay()
        return self.out


class GraphedLogProb:
    """`model.log_prob(x)` without gradients (evaluation; core.py:170-183) captured once for a fixed batch shape and
    replayed: `lp = GraphedLogProb(model, x_example)(x)` -- x is copied into the static input, the result is a static
    tensor overwritten by the next call.  Falls back to eager calls if capture is not possible."""

    def __init__(self, model, x_example, warmup=3):
        self.model = model
        self.x = x_example.detach().clone()
        self.graph, self.out, self.capture_error = None, None, None
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side), torch.no_grad():
            for _ in range(warmup):
                model.log_prob(self.x)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        try:
            g = torch.cuda.CUDAGraph()
            with torch.no_grad(), torch.cuda.graph(g):
                self.out = model.log_prob(self.x)
            self.graph = g
        except Exception as exc:            # noqa: BLE001 -- report and run eagerly
            self.capture_error = _after_failed_capture(exc)
            self.graph = None

    def __call__(self, x):
        if self.graph is None:
            with torch.no_grad():
                return self.model.log_prob(x)
        self.x.copy_(x, non_blocking=True)
        self.graph.replay()
        return self.out
