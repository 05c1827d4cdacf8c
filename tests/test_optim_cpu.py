"""Host-side logic of the flat optimiser path that needs no GPU: bucket layout and the loud failure on CPU
parameters (the product has no CPU path)."""
import pytest
import torch

from vcnf_b200 import optim, parallel


def test_grad_bucket_aligned_layout_and_collect():
    ps = [torch.nn.Parameter(torch.randn(n)) for n in (5, 33, 64, 1)]
    b = parallel.GradBucket(ps, align=32)
    assert b.offsets == [0, 32, 96, 160] and b.total == 192
    assert all(p.grad.data_ptr() == b.flat[o:].data_ptr() for p, o in zip(ps, b.offsets))
    b.begin()
    assert all(p.grad is None for p in ps)
    fresh = [torch.full_like(p, float(i + 1)) for i, p in enumerate(ps)]
    for p, g in zip(ps[:3], fresh[:3]):
        p.grad = g                                   # the last parameter received no gradient
    b.collect()
    for i, (p, o) in enumerate(zip(ps, b.offsets)):
        assert p.grad.data_ptr() == b.flat[o:].data_ptr()
        want = float(i + 1) if i < 3 else 0.0
        assert torch.all(b.flat[o:o + p.numel()] == want)
    pad = torch.ones(192, dtype=torch.bool)
    for p, o in zip(ps, b.offsets):
        pad[o:o + p.numel()] = False
    assert torch.all(b.flat[pad] == 0)               # pad elements stay zero: Adam leaves them at zero


def test_grad_bucket_default_layout_is_dense():
    ps = [torch.nn.Parameter(torch.randn(n)) for n in (5, 33)]
    b = parallel.GradBucket(ps)
    assert b.offsets == [0, 5] and b.total == 38


def test_flat_adam_refuses_cpu_parameters():
    with pytest.raises(RuntimeError):
        optim.FlatAdam([torch.nn.Parameter(torch.randn(4))])
    with pytest.raises(ValueError):
        optim.FlatAdam([])
