"""Base distributions at the two ends of the flow (reference: normflow/distributions/base.py).  On a CUDA device
DiagGaussian / ClassCondDiagGaussian (log_prob and the reparametrised draw) and GaussianMixture.log_prob are one
kernel each with hand-written backward (vcnf_gauss_fwd / _bwd, vcnf_gmm_logprob / _bwd); the
tensor-op bodies below are the host mirror the CPU tests and soft-label calls use."""
import numpy as np
import torch
from torch import nn

from .. import ops


class BaseDistribution(nn.Module):
    def forward(self, num_samples=1):
        raise NotImplementedError

    def log_prob(self, z):
        raise NotImplementedError


class DiagGaussian(BaseDistribution):
    """base.py:609-652."""

    def __init__(self, shape, trainable=True):
        super().__init__()
        if isinstance(shape, int):
            shape = (shape,)
        self.shape = tuple(shape)
        self.n_dim = len(self.shape)
        self.d = np.prod(self.shape)
        if trainable:
            self.loc = nn.Parameter(torch.zeros(1, *self.shape))
            self.log_scale = nn.Parameter(torch.zeros(1, *self.shape))
        else:
            self.register_buffer("loc", torch.zeros(1, *self.shape))
            self.register_buffer("log_scale", torch.zeros(1, *self.shape))
        self.temperature = None

    def _log_scale(self):
        if self.temperature is None:
            return self.log_scale
        return self.log_scale + np.log(self.temperature)

    def _log_temp(self):
        return 0.0 if self.temperature is None else float(np.log(self.temperature))

    def forward(self, num_samples=1):
        eps = torch.randn((num_samples,) + self.shape, dtype=self.loc.dtype, device=self.loc.device)
        if eps.is_cuda:
            return ops.gauss_sample(eps, self.loc, self.log_scale, log_temp=self._log_temp())
        log_scale = self._log_scale()
        z = self.loc + torch.exp(log_scale) * eps
        log_p = - 0.5 * self.d * np.log(2 * np.pi) \
            - torch.sum(log_scale + 0.5 * torch.pow(eps, 2), list(range(1, self.n_dim + 1)))
        return z, log_p

    def log_prob(self, z):
        if z.is_cuda:
            return ops.gauss_logprob(z, self.loc, self.log_scale, log_temp=self._log_temp())
        log_scale = self._log_scale()
        return - 0.5 * self.d * np.log(2 * np.pi) \
            - torch.sum(log_scale + 0.5 * torch.pow((z - self.loc) / torch.exp(log_scale), 2),
                        list(range(1, self.n_dim + 1)))


class ClassCondDiagGaussian(BaseDistribution):
    """base.py:715-775."""

    def __init__(self, shape, num_classes):
        super().__init__()
        if isinstance(shape, int):
            shape = (shape,)
        self.shape = tuple(shape)
        self.n_dim = len(self.shape)
        self.perm = [self.n_dim] + list(range(self.n_dim))
        self.d = np.prod(self.shape)
        self.num_classes = num_classes
        self.loc = nn.Parameter(torch.zeros(*self.shape, num_classes))
        self.log_scale = nn.Parameter(torch.zeros(*self.shape, num_classes))
        self.temperature = None

    def _class_params(self, y):
        if y.dim() == 1:
            onehot = torch.zeros((self.num_classes, len(y)), dtype=self.loc.dtype, device=self.loc.device)
            onehot.scatter_(0, y[None], 1)
            y = onehot
        else:
            y = y.t()
        loc = (self.loc @ y).permute(*self.perm)
        log_scale = (self.log_scale @ y).permute(*self.perm)
        if self.temperature is not None:
            log_scale = np.log(self.temperature) + log_scale
        return loc, log_scale

    def forward(self, num_samples=1, y=None):
        if y is not None:
            num_samples = len(y)
        else:
            y = torch.randint(self.num_classes, (num_samples,), device=self.loc.device)
        eps = torch.randn((num_samples,) + self.shape, dtype=self.loc.dtype, device=self.loc.device)
        if eps.is_cuda and y.dim() == 1:
            return ops.gauss_sample(eps, self.loc, self.log_scale, y=y.long(), classes=self.num_classes,
                                    log_temp=0.0 if self.temperature is None else float(np.log(self.temperature)))
        loc, log_scale = self._class_params(y)
        z = loc + torch.exp(log_scale) * eps
        log_p = - 0.5 * self.d * np.log(2 * np.pi) \
            - torch.sum(log_scale + 0.5 * torch.pow(eps, 2), list(range(1, self.n_dim + 1)))
        return z, log_p

    def log_prob(self, z, y):
        if z.is_cuda and y.dim() == 1:
            return ops.gauss_logprob(z, self.loc, self.log_scale, y=y.long(), classes=self.num_classes,
                                     log_temp=0.0 if self.temperature is None else float(np.log(self.temperature)))
        loc, log_scale = self._class_params(y)
        return - 0.5 * self.d * np.log(2 * np.pi) \
            - torch.sum(log_scale + 0.5 * torch.pow((z - loc) / torch.exp(log_scale), 2),
                        list(range(1, self.n_dim + 1)))


class GaussianMixture(BaseDistribution):
    """base.py:957-1034."""

    def __init__(self, n_modes, dim, loc=None, scale=None, weights=None, trainable=True):
        super().__init__()
        self.n_modes = n_modes
        self.dim = dim
        loc = np.zeros((n_modes, dim)) if loc is None else np.array(loc)
        scale = np.ones((n_modes, dim)) if scale is None else np.array(scale)
        weights = (np.ones(n_modes) + 1.) if weights is None else np.array(weights, dtype=float)
        loc, scale, weights = loc[None, ...], scale[None, ...], weights[None, ...]
        weights = weights / weights.sum(1)
        tensors = {"loc": torch.tensor(1. * loc), "log_scale": torch.tensor(np.log(1. * scale)),
                   "weight_scores": torch.tensor(np.log(1. * weights))}
        for name, value in tensors.items():
            if trainable:
                setattr(self, name, nn.Parameter(value))
            else:
                self.register_buffer(name, value)

    def log_prob(self, z):
        if z.is_cuda and z.dim() == 2:
            return ops.gmm_logprob(z, self.loc, self.log_scale, self.weight_scores)
        weights = torch.softmax(self.weight_scores, 1)
        eps = (z[:, None, :] - self.loc) / torch.exp(self.log_scale)
        log_p = - 0.5 * self.dim * np.log(2 * np.pi) + torch.log(weights) \
            - 0.5 * torch.sum(torch.pow(eps, 2), 2) - torch.sum(self.log_scale, 2)
        return torch.logsumexp(log_p, 1)

    def forward(self, num_samples=1):
        weights = torch.softmax(self.weight_scores, 1)
        mode = torch.multinomial(weights[0, :], num_samples, replacement=True)
        mode_1h = nn.functional.one_hot(mode, self.n_modes)[..., None]
        eps_ = torch.randn(num_samples, self.dim, dtype=self.loc.dtype, device=self.loc.device)
        scale_sample = torch.sum(torch.exp(self.log_scale) * mode_1h, 1)
        loc_sample = torch.sum(self.loc * mode_1h, 1)
        z = eps_ * scale_sample + loc_sample
        return z, self.log_prob(z)
