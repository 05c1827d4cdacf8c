"""Optimiser step of the training loops on flat buffers.

The reference trains with `torch.optim.Adam(nfm.parameters(), lr=1e-4, weight_decay=1e-6)` (runcred.py:192,
run.py:139) and calls `optimizer.step()` after `loss.backward()` (runcred.py:246-258).  A cfg-2 model has ~260
parameter tensors; the library's multi-tensor Adam needs 0.58 ms per step for 1.8 M parameters (chunk tables, one
step counter per tensor).  `FlatAdam` re-homes the parameters into ONE buffer (each `p.data` becomes a view of it),
keeps both moments flat, takes the gradients from a `parallel.GradBucket` laid out the same way and runs ONE
kernel (`vcnf_adam_flat`, 28 B per fp32 parameter).  Arithmetic and hyper-parameters are torch.optim.Adam's
(amsgrad off, maximize off); a parameter that received no gradient counts as a zero gradient.
Build it AFTER the model is on its device and in its dtype (`nfm.double().to(device)` first, like the run scripts
do before creating their optimiser): `.to()` / `.double()` afterwards re-allocates the parameters and cuts
them loose from the flat buffer."""
import torch

from . import ops
from ._lib import ptr, stream, dtype_code
from .parallel import GradBucket

ALIGN = 32          # elements: every parameter starts on a 128 B (fp32) / 256 B (fp64) boundary


class FlatAdam(torch.optim.Optimizer):
    _entry = "vcnf_adam_flat_guarded"

    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, bucket=None):
        params = [p for p in params if p.requires_grad]
        if not params:
            raise ValueError("FlatAdam: no parameters to optimise")
        if not all(p.is_cuda for p in params):
            raise RuntimeError("FlatAdam runs on CUDA parameters only (vcnf_b200 has no CPU path)")
        if lr < 0 or eps < 0 or weight_decay < 0 or not (0 <= betas[0] < 1 and 0 <= betas[1] < 1):
            raise ValueError("FlatAdam: invalid hyper-parameter")
        super().__init__(params, dict(lr=lr, betas=tuple(betas), eps=eps, weight_decay=weight_decay))
        if bucket is None:
            bucket = GradBucket(params, align=ALIGN)
        if [id(p) for p in bucket.params] != [id(p) for p in params] or any(o % ALIGN for o in bucket.offsets):
            raise ValueError("FlatAdam: the gradient bucket must hold the same parameters in the same order, "
                             "GradBucket(params, align=optim.ALIGN)")
        self.bucket = bucket
        p0 = params[0]
        self.flat_p = torch.zeros(bucket.total, dtype=p0.dtype, device=p0.device)
        with torch.no_grad():
            for p, off in zip(params, bucket.offsets):
                view = self.flat_p[off:off + p.numel()].view_as(p)
                view.copy_(p)
                p.data = view                 # the module keeps its Parameter objects; only the storage moves
        self.exp_avg = torch.zeros_like(self.flat_p)
        self.exp_avg_sq = torch.zeros_like(self.flat_p)
        self.steps = torch.zeros(1, dtype=torch.float32, device=p0.device)
        self._ticket = torch.zeros(1, dtype=torch.int32, device=p0.device)
        # the learning rate lives in a device scalar the kernel reads, so a step replayed from a CUDA graph follows
        # later edits of param_groups[0]["lr"] (sync_hyper() is called by step() and by GraphedTrainStep)
        self._lr_dev = torch.full((1,), float(lr), dtype=torch.float64, device=p0.device)
        self._lr_host = float(lr)
        self.guard = None          # device scalar (the loss): a non-finite value skips the update (runcred.py:249-250)
        self.grad_scale = 1.0      # folded into the kernel: 1 / world_size after a summing all-reduce
        self._skip = torch.full((1,), float("nan"), dtype=torch.float32, device=p0.device)
        self.frozen = False        # True: every step is skipped (GraphedTrainStep's warm-up iterations)

    def add_param_group(self, group):
        if getattr(self, "param_groups", None):
            raise ValueError("FlatAdam keeps ONE parameter group (one flat buffer, one set of hyper-parameters)")
        super().add_param_group(group)

    def sync_hyper(self):
        """Push a changed learning rate to the device scalar (outside graph capture)."""
        lr = float(self.param_groups[0]["lr"])
        if lr != self._lr_host and not torch.cuda.is_current_stream_capturing():
            self._lr_dev.fill_(lr)
            self._lr_host = lr

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        g = self.param_groups[0]
        self.sync_hyper()
        self.bucket.collect()                 # no-op when the gradients already are the bucket's views
        guard = self._skip if self.frozen else self.guard
        if guard is not None and guard.dtype not in (torch.float32, torch.float64):
            guard = guard.float()
        ops._launch(self._entry, ptr(self.flat_p), ptr(self.bucket.flat), ptr(self.exp_avg),
                    ptr(self.exp_avg_sq), ptr(self.steps), ptr(self._ticket), self.bucket.total, ptr(self._lr_dev),
                    float(g["betas"][0]), float(g["betas"][1]), float(g["eps"]), float(g["weight_decay"]),
                    float(self.grad_scale), ptr(guard), dtype_code(guard) if guard is not None else 0,
                    dtype_code(self.flat_p), stream())
        return loss

    def zero_grad(self, set_to_none=True):
        if set_to_none:
            self.bucket.begin()
        else:
            self.bucket.collect()
            self.bucket.zero()

    def state_dict(self):
        return {"exp_avg": self.exp_avg.clone(), "exp_avg_sq": self.exp_avg_sq.clone(), "steps": self.steps.clone(),
                "param_groups": [{k: v for k, v in g.items() if k != "params"} for g in self.param_groups]}

    def load_state_dict(self, state):
        self.exp_avg.copy_(state["exp_avg"])
        self.exp_avg_sq.copy_(state["exp_avg_sq"])
        self.steps.copy_(state["steps"])
        for g, saved in zip(self.param_groups, state["param_groups"]):
            g.update(saved)


class FlatAdamax(FlatAdam):
    """torch.optim.Adamax (the Glow notebook's optimiser, example/glow.ipynb:88: lr 1e-3, weight_decay 1e-5) on the same
    flat buffers, one launch (`vcnf_adamax_flat_guarded`): `exp_avg_sq` holds the running infinity norm
    u = max(beta2 u, |g| + eps); p -= lr / (1 - beta1^t) m / u.  The library's capturable Adamax is ~70 multi-tensor
    launches per step on the 1 400 tensors of cfg 5."""
    _entry = "vcnf_adamax_flat_guarded"

    def __init__(self, params, lr=2e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, bucket=None):
        super().__init__(params, lr=lr, betas=betas, eps=eps, weight_decay=weight_decay, bucket=bucket)
