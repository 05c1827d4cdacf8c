"""Base distributions the reference's run scripts choose between besides the three of SURVEY row f1
(runcred.py:92-117, run.py:73-97: `nf.distributions.base.MultivariateGaussian / T / GGD / GMM`).

They sit at the latent end of `forward_kld` / `sample` exactly like DiagGaussian, but they are a few `torch.distributions`
calls on [B, D] and none of them is part of a BASELINE configuration, so they run as library ops on whatever device their
parameters live on (the reference hard-codes 'cuda'; here: CUDA when present).  What is mirrored is the reference's
BEHAVIOUR, quirks included, so that a script pointed at this package computes what it computed before:

* `T` and `GGD` return the MEAN over features of the per-feature log-densities (base.py:507,511,546,550), not the sum;
* `GGD`'s density is the Gaussian one -- the exponent p only shapes the samples (base.py:142-148);
* a non-trainable `T` has df + 1 degrees of freedom, a trainable one df (base.py:531-533);
* `MultivariateGaussian`'s covariance is identity + scale in EVERY entry (base.py:306,310), and the
  `torch.distributions` object is built once in the constructor, i.e. it keeps the initial parameters (base.py:312);
* `GMM.log_prob` evaluates the mixture built by the last `forward` call (or the constructor) (base.py:202-207,233).
"""
import math

import torch
from torch import nn
from torch import distributions as D

from .base import BaseDistribution


def _device():
    return torch.device("cuda" if torch.cuda.is_available() else "cpu")


def _param_or_buffer(module, name, value, trainable):
    if trainable:
        setattr(module, name, nn.Parameter(value, requires_grad=True))
    else:
        module.register_buffer(name, value)


class MultivariateGaussian(BaseDistribution):
    """base.py:288-325."""

    def __init__(self, n_dim=2, n_components=3, trainable=False, loc=0., scale=1.0):
        super().__init__()
        self.n_components, self.n_dim = n_components, n_dim
        dev = _device()
        _param_or_buffer(self, "loc", torch.zeros(n_dim, dtype=torch.double, device=dev) + loc, trainable)
        _param_or_buffer(self, "scale", torch.eye(n_dim, dtype=torch.double, device=dev) + scale, trainable)
        self.mvn = D.MultivariateNormal(self.loc, self.scale)

    def forward(self, num_samples=1):
        z = self.mvn.sample([num_samples])
        return z, self.mvn.log_prob(z)

    def log_prob(self, z):
        return self.mvn.log_prob(z)


class T(BaseDistribution):
    """Independent Student-t features, mean log-density over the features (base.py:513-550)."""

    def __init__(self, n_dim=2, df=2., trainable=False):
        super().__init__()
        self.n_dim = n_dim
        dev = _device()
        start = torch.zeros if trainable else torch.ones
        _param_or_buffer(self, "df", start((n_dim,), dtype=torch.double, device=dev) + df, trainable)
        self.t = D.StudentT(self.df)

    def forward(self, num_samples=1):
        z = self.t.sample([num_samples])
        return z, self.t.log_prob(z).mean(axis=1)

    def log_prob(self, z):
        return self.t.log_prob(z).mean(axis=1)


class GenNormal:
    """What the reference's `GenNormal` (base.py:28-160) computes: samples loc + scale * s * G^(1/p) with G ~ Gamma(1/p, 1)
    and a random sign s, and a log-density that is the Gaussian one in (loc, scale)."""

    def __init__(self, loc, scale, p):
        self.loc, self.scale, self.p = loc, scale, p

    def sample(self, sample_shape=torch.Size()):
        shape = torch.Size(sample_shape) + self.loc.shape
        with torch.no_grad():
            ipower = 1.0 / self.p
            g = D.Gamma(ipower, torch.ones_like(ipower)).sample(torch.Size(sample_shape))
            sign = torch.randint(0, 2, shape, device=g.device).to(self.loc.dtype) * 2 - 1
            return self.loc + self.scale * sign * torch.pow(g.abs(), ipower)

    def log_prob(self, value):
        var = self.scale ** 2
        return -((value - self.loc) ** 2) / (2 * var) - self.scale.log() - math.log(math.sqrt(2 * math.pi))


class GGD(BaseDistribution):
    """base.py:469-511."""

    def __init__(self, n_dim=2, beta=2., trainable=False):
        super().__init__()
        self.n_dim, self.beta = n_dim, beta
        dev = _device()
        _param_or_buffer(self, "p", torch.ones((n_dim,), dtype=torch.double, device=dev) + beta, trainable)
        _param_or_buffer(self, "loc", torch.zeros(n_dim, dtype=torch.double, device=dev), trainable)
        _param_or_buffer(self, "scale", torch.ones(n_dim, dtype=torch.double, device=dev), trainable)
        self.ggd = GenNormal(self.loc, self.scale, self.p)

    def forward(self, num_samples=1):
        z = self.ggd.sample([num_samples])
        return z, self.ggd.log_prob(z).mean(axis=1)

    def log_prob(self, z):
        return self.ggd.log_prob(z).mean(axis=1)


class GMM(nn.Module):
    """Mixture of n_cell isotropic Gaussians whose means / widths lie on logarithmic grids with learnable bases
    (base.py:188-234): mean_k = log(scale k + shift) / log(mbase), std_k = log(scale k + shift) / log(vbase) + 1e-3."""

    def __init__(self, weights, mbase, vbase, scale, n_cell=8, shift=0, dim=2, trainable=False):
        super().__init__()
        self.weight = nn.Parameter(weights, requires_grad=True)
        self.mbase = nn.Parameter(mbase, requires_grad=True)
        self.vbase = nn.Parameter(vbase, requires_grad=True)
        self.scale = nn.Parameter(scale, requires_grad=False)
        self.grid = torch.arange(1, n_cell + 1, device=weights.device)
        self.shift, self.n_cell, self.dim = shift, n_cell, dim
        with torch.no_grad():
            self.gmm = self._mixture()

    def _grid(self, base):
        return (torch.log(self.scale * self.grid + self.shift) / torch.log(base)).reshape(-1, 1).expand(self.n_cell, self.dim)

    def trsf_gridm(self):
        return self._grid(self.mbase)

    def trsf_gridv(self):
        return self._grid(self.vbase)

    def _mixture(self):
        comp = D.Independent(D.Normal(self.trsf_gridm(), self.trsf_gridv() + 0.001), 1)
        return D.MixtureSameFamily(D.Categorical(self.weight), comp)

    def forward(self, num_samples=1):
        self.gmm = self._mixture()
        with torch.no_grad():
            samples = self.gmm.sample([num_samples]).double()
        return samples, self.gmm.log_prob(samples)

    def log_prob(self, z):
        return self.gmm.log_prob(z)
