"""Small helpers (reference: normflow/utils/nn.py)."""
import torch
from torch import nn


class ConstScaleLayer(nn.Module):
    def __init__(self, scale=1.):
        super().__init__()
        self.scale_cpu = torch.tensor(scale)
        self.register_buffer("scale", self.scale_cpu)

    def forward(self, input):
        return input * self.scale


class ClampExp(nn.Module):
    def forward(self, x):
        one = torch.tensor(1., device=x.device, dtype=x.dtype)
        return torch.min(torch.exp(x), one)


class PeriodicFeatures(nn.Module):
    """w0*sin(scale*x) + w1*cos(scale*x) on the circular coordinates (nn.py:59-119); part of
    the conditioner input pipeline of the circular spline layers."""

    def __init__(self, ndim, ind, scale=1., bias=False, activation=None):
        super().__init__()
        self.ndim = ndim
        if torch.is_tensor(ind):
            self.register_buffer("ind", ind.long())
        else:
            self.register_buffer("ind", torch.tensor(ind, dtype=torch.long))
        ind_ = [i for i in range(self.ndim) if i not in self.ind]
        self.register_buffer("ind_", torch.tensor(ind_, dtype=torch.long))
        perm_ = torch.cat((self.ind, self.ind_))
        inv_perm_ = torch.zeros_like(perm_)
        for i in range(self.ndim):
            inv_perm_[perm_[i]] = i
        self.register_buffer("inv_perm", inv_perm_)
        self.weights = nn.Parameter(torch.ones(len(self.ind), 2))
        if torch.is_tensor(scale):
            self.register_buffer("scale", scale)
        else:
            self.scale = scale
        self.apply_bias = bias
        if self.apply_bias:
            self.bias = nn.Parameter(torch.zeros(len(self.ind)))
        self.activation = (lambda x: x) if activation is None else activation

    def forward(self, inputs):
        inputs_ = inputs[..., self.ind]
        inputs_ = self.scale * inputs_
        inputs_ = self.weights[:, 0] * torch.sin(inputs_) + self.weights[:, 1] * torch.cos(inputs_)
        if self.apply_bias:
            inputs_ = inputs_ + self.bias
        inputs_ = self.activation(inputs_)
        out = torch.cat((inputs_, inputs[..., self.ind_]), -1)
        return out[..., self.inv_perm]


def tile(x, n):
    x_ = x.reshape(-1)
    x_ = x_.repeat(n)
    x_ = x_.reshape(n, -1)
    x_ = x_.transpose(1, 0)
    return x_.reshape(-1)


def sum_except_batch(x, num_batch_dims=1):
    """Host-side helper kept for API parity (nn.py:131-134); inside the layers this reduction is
    fused into the kernels."""
    return torch.sum(x, dim=list(range(num_batch_dims, x.ndimension())))
