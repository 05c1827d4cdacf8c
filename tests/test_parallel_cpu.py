"""World-size-2 `gloo` tests of the data-parallel plumbing (vcnf_b200/parallel.py) on CPU: batch
shards, the flat gradient bucket and its single all-reduce.  The bijector kernels themselves need a
GPU; the host logic does not, so a small torch module stands in for the flow here."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from vcnf_b200 import parallel


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _model():
    torch.manual_seed(7)
    return torch.nn.Sequential(torch.nn.Linear(6, 16), torch.nn.Tanh(), torch.nn.Linear(16, 1)).double()


def _loss(model, x):
    return model(x).square().mean()


def _worker(rank, world, port, n, out):
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank),
                      MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    r, w, _ = parallel.init_from_env(backend="gloo")
    assert (r, w) == (rank, world)
    torch.manual_seed(0)
    x = torch.randn(n, 6, dtype=torch.float64)
    lo, hi = parallel.shard_bounds(n, rank, world)
    model = _model()
    bucket = parallel.GradBucket(model.parameters())
    bucket.zero()
    # weight each shard by its share of the batch so that ragged shards average exactly
    (_loss(model, x[lo:hi]) * ((hi - lo) * world / n)).backward()
    bucket.allreduce_mean()
    if rank == 0:
        torch.save({"flat": bucket.flat.clone(), "bounds": (lo, hi)}, out)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("n", [64, 37])
def test_two_rank_bucket_allreduce_matches_full_batch(tmp_path, n):
    out = str(tmp_path / "r0.pt")
    mp.spawn(_worker, args=(2, _free_port(), n, out), nprocs=2, join=True)
    got = torch.load(out)
    torch.manual_seed(0)
    x = torch.randn(n, 6, dtype=torch.float64)
    model = _model()
    _loss(model, x).backward()
    want = torch.cat([p.grad.reshape(-1) for p in model.parameters()])
    assert got["bounds"] == (0, (n + 1) // 2)
    torch.testing.assert_close(got["flat"], want, rtol=1e-12, atol=1e-14)


@pytest.mark.parametrize("n,world", [(0, 2), (1, 2), (7, 4), (65536, 8), (262144, 8), (10, 3)])
def test_shard_bounds_partition_the_batch(n, world):
    spans = [parallel.shard_bounds(n, r, world) for r in range(world)]
    assert spans[0][0] == 0 and spans[-1][1] == n
    for a, b in zip(spans, spans[1:]):
        assert a[1] == b[0]
    sizes = [hi - lo for lo, hi in spans]
    assert max(sizes) - min(sizes) <= 1


def test_bucket_grads_are_views_of_one_buffer():
    model = _model()
    bucket = parallel.GradBucket(model.parameters())
    _loss(model, torch.randn(8, 6, dtype=torch.float64)).backward()
    base = bucket.flat.untyped_storage().data_ptr()
    assert all(p.grad.untyped_storage().data_ptr() == base for p in model.parameters())
    assert bucket.flat.abs().sum() > 0
    bucket.zero()
    assert all(float(p.grad.abs().sum()) == 0.0 for p in model.parameters())


def _worker_collect(rank, world, port, n, out):
    os.environ.update(RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank),
                      MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    parallel.init_from_env(backend="gloo")
    torch.manual_seed(0)
    x = torch.randn(n, 6, dtype=torch.float64)
    lo, hi = parallel.shard_bounds(n, rank, world)
    model = _model()
    bucket = parallel.GradBucket(model.parameters())
    for _ in range(2):                         # second round: .grad starts as the bound views again
        bucket.begin()
        assert all(p.grad is None for p in model.parameters())
        (_loss(model, x[lo:hi]) * ((hi - lo) * world / n)).backward()
        bucket.collect()
        bucket.allreduce_mean()
    base = bucket.flat.untyped_storage().data_ptr()
    assert all(p.grad.untyped_storage().data_ptr() == base for p in model.parameters())
    if rank == 0:
        torch.save({"flat": bucket.flat.clone()}, out)
    dist.barrier()
    dist.destroy_process_group()


def test_begin_collect_protocol_two_ranks(tmp_path):
    """begin() / collect(): gradients are produced as fresh tensors (no per-parameter accumulate launches),
    packed into the flat buffer by one multi-tensor copy, then all-reduced -- same result as the full batch."""
    out = str(tmp_path / "r0.pt")
    n = 50
    mp.spawn(_worker_collect, args=(2, _free_port(), n, out), nprocs=2, join=True)
    got = torch.load(out)
    torch.manual_seed(0)
    x = torch.randn(n, 6, dtype=torch.float64)
    model = _model()
    _loss(model, x).backward()
    want = torch.cat([p.grad.reshape(-1) for p in model.parameters()])
    torch.testing.assert_close(got["flat"], want, rtol=1e-12, atol=1e-14)


def test_collect_zero_fills_parameters_without_gradient():
    model = _model()
    extra = torch.nn.Parameter(torch.ones(3, dtype=torch.float64))
    bucket = parallel.GradBucket(list(model.parameters()) + [extra])
    bucket.flat.fill_(7.0)
    bucket.begin()
    _loss(model, torch.randn(8, 6, dtype=torch.float64)).backward()
    bucket.collect()
    assert float(extra.grad.abs().sum()) == 0.0
    assert bucket.flat[:-3].abs().sum() > 0
