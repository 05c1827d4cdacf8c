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

    def __init__(self, params):
        self.params = [p for p in params if p.requires_grad]
        if not self.params:
            self.flat = None
            return
        p0 = self.params[0]
        total = sum(p.numel() for p in self.params)
        self.flat = torch.zeros(total, dtype=p0.dtype, device=p0.device)
        off = 0
        for p in self.params:
            if p.dtype != p0.dtype:
                raise TypeError("GradBucket needs a single parameter dtype")
            p.grad = self.flat[off:off + p.numel()].view_as(p)
            off += p.numel()

    def zero(self):
        if self.flat is not None:
            self.flat.zero_()

    def allreduce_mean(self, world=None):
        if self.flat is None or not dist.is_initialized():
            return
        world = dist.get_world_size() if world is None else world
        if world > 1:
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM)
            self.flat.div_(world)


def sync_actnorm_init(model, src=0):
    """After the first (initialising) step make every rank use rank `src`'s data-dependent
    ActNorm statistics (flows/normalization.py:22-27 is per-process in the reference)."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return
    for m in model.modules():
        if hasattr(m, "data_dep_init_done") and hasattr(m, "s") and hasattr(m, "t"):
            dist.broadcast(m.s.data, src)
            dist.broadcast(m.t.data, src)
