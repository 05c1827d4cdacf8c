"""Oracle base distributions / targets at the two ends of the flow (plain torch).

Only the ones the BASELINE configs use: `DiagGaussian` (distributions/base.py:609-652),
`ClassCondDiagGaussian` (:715-775), `GaussianMixture` (:957-1034), `TwoMoons`
(distributions/target.py:275-296).
"""
import math

import numpy as np
import torch
from torch import nn

_LOG_2PI = math.log(2 * math.pi)


class DiagGaussian(nn.Module):
    def __init__(self, shape, trainable=True):
        super().__init__()
        shape = (shape,) if isinstance(shape, int) else tuple(shape)
        self.shape, self.n_dim, self.d = shape, len(shape), int(np.prod(shape))
        for name in ("loc", "log_scale"):
            t = torch.zeros(1, *shape)
            if trainable:
                setattr(self, name, nn.Parameter(t))
            else:
                self.register_buffer(name, t)
        self.temperature = None

    def _log_scale(self):
        return self.log_scale if self.temperature is None else self.log_scale + math.log(self.temperature)

    def forward(self, num_samples=1):
        eps = torch.randn((num_samples,) + self.shape, dtype=self.loc.dtype, device=self.loc.device)
        ls = self._log_scale()
        z = self.loc + torch.exp(ls) * eps
        dims = list(range(1, self.n_dim + 1))
        return z, -0.5 * self.d * _LOG_2PI - torch.sum(ls + 0.5 * torch.pow(eps, 2), dims)

    def log_prob(self, z):
        ls = self._log_scale()
        dims = list(range(1, self.n_dim + 1))
        return -0.5 * self.d * _LOG_2PI - torch.sum(
            ls + 0.5 * torch.pow((z - self.loc) / torch.exp(ls), 2), dims)


class ClassCondDiagGaussian(nn.Module):
    def __init__(self, shape, num_classes):
        super().__init__()
        shape = (shape,) if isinstance(shape, int) else tuple(shape)
        self.shape, self.n_dim, self.d = shape, len(shape), int(np.prod(shape))
        self.perm = [self.n_dim] + list(range(self.n_dim))
        self.num_classes = num_classes
        self.loc = nn.Parameter(torch.zeros(*shape, num_classes))
        self.log_scale = nn.Parameter(torch.zeros(*shape, num_classes))
        self.temperature = None

    def _onehot(self, y):
        if y.dim() != 1:
            return y.t()
        oh = torch.zeros((self.num_classes, len(y)), dtype=self.loc.dtype, device=self.loc.device)
        oh.scatter_(0, y[None], 1)
        return oh

    def _params(self, y):
        loc = (self.loc @ y).permute(*self.perm)
        ls = (self.log_scale @ y).permute(*self.perm)
        if self.temperature is not None:
            ls = math.log(self.temperature) + ls
        return loc, ls

    def forward(self, num_samples=1, y=None):
        if y is None:
            y = torch.randint(self.num_classes, (num_samples,), device=self.loc.device)
        else:
            num_samples = len(y)
        loc, ls = self._params(self._onehot(y))
        eps = torch.randn((num_samples,) + self.shape, dtype=self.loc.dtype, device=self.loc.device)
        dims = list(range(1, self.n_dim + 1))
        return loc + torch.exp(ls) * eps, \
            -0.5 * self.d * _LOG_2PI - torch.sum(ls + 0.5 * torch.pow(eps, 2), dims)

    def log_prob(self, z, y):
        loc, ls = self._params(self._onehot(y))
        dims = list(range(1, self.n_dim + 1))
        return -0.5 * self.d * _LOG_2PI - torch.sum(
            ls + 0.5 * torch.pow((z - loc) / torch.exp(ls), 2), dims)


class GaussianMixture(nn.Module):
    def __init__(self, n_modes, dim, loc=None, scale=None, weights=None, trainable=True):
        super().__init__()
        self.n_modes, self.dim = n_modes, dim
        loc = np.zeros((n_modes, dim)) if loc is None else np.array(loc)
        scale = np.ones((n_modes, dim)) if scale is None else np.array(scale)
        weights = np.ones(n_modes) + 1.0 if weights is None else np.array(weights, dtype=float)
        weights = weights[None] / weights.sum()
        vals = {"loc": torch.tensor(1.0 * loc[None]), "log_scale": torch.tensor(np.log(1.0 * scale[None])),
                "weight_scores": torch.tensor(np.log(weights))}
        for k, v in vals.items():
            if trainable:
                setattr(self, k, nn.Parameter(v))
            else:
                self.register_buffer(k, v)

    def log_prob(self, z):
        w = torch.softmax(self.weight_scores, 1)
        eps = (z[:, None, :] - self.loc) / torch.exp(self.log_scale)
        lp = -0.5 * self.dim * _LOG_2PI + torch.log(w) - 0.5 * torch.sum(torch.pow(eps, 2), 2) \
            - torch.sum(self.log_scale, 2)
        return torch.logsumexp(lp, 1)

    def forward(self, num_samples=1):
        w = torch.softmax(self.weight_scores, 1)
        mode = torch.multinomial(w[0, :], num_samples, replacement=True)
        oh = nn.functional.one_hot(mode, self.n_modes)[..., None]
        eps = torch.randn(num_samples, self.dim, dtype=self.loc.dtype, device=self.loc.device)
        z = eps * torch.sum(torch.exp(self.log_scale) * oh, 1) + torch.sum(self.loc * oh, 1)
        return z, self.log_prob(z)


class TwoMoons(nn.Module):
    def __init__(self):
        super().__init__()
        self.n_dims = 2
        # `Target.__init__` (distributions/target.py:14-23) registers the proposal box as buffers
        self.register_buffer("prop_scale", torch.tensor(20.0))
        self.register_buffer("prop_shift", torch.tensor(-10.0))

    def log_prob(self, z):
        a = torch.abs(z[:, 0])
        return -0.5 * ((torch.norm(z, dim=1) - 2) / 0.2) ** 2 - 0.5 * ((a - 2) / 0.3) ** 2 \
            + torch.log(1 + torch.exp(-4 * a / 0.09))
