"""Target densities used to drive the configs (reference: normflow/distributions/target.py)."""
import numpy as np
import torch
from torch import nn


class Target(nn.Module):
    """target.py:10-62: rejection sampling from a uniform proposal box."""

    def __init__(self, prop_scale=torch.tensor(20.), prop_shift=torch.tensor(-10.)):
        super().__init__()
        self.register_buffer("prop_scale", prop_scale)
        self.register_buffer("prop_shift", prop_shift)

    def log_prob(self, z):
        raise NotImplementedError("The log probability is not implemented yet.")

    def rejection_sampling(self, num_steps=1):
        eps = torch.rand((num_steps, self.n_dims), dtype=self.prop_scale.dtype, device=self.prop_scale.device)
        z_ = self.prop_scale * eps + self.prop_shift
        prob = torch.rand(num_steps, dtype=self.prop_scale.dtype, device=self.prop_scale.device)
        prob_ = torch.exp(self.log_prob(z_) - self.max_log_prob)
        return z_[prob_ > prob, :]

    def sample(self, num_samples=1):
        z = torch.zeros((0, self.n_dims), dtype=self.prop_scale.dtype, device=self.prop_scale.device)
        while len(z) < num_samples:
            z_ = self.rejection_sampling(num_samples)
            ind = np.min([len(z_), num_samples - len(z)])
            z = torch.cat([z, z_[:ind, :]], 0)
        return z


class TwoMoons(Target):
    """target.py:275-296."""

    def __init__(self):
        super().__init__()
        self.n_dims = 2
        self.max_log_prob = 0.

    def log_prob(self, z):
        a = torch.abs(z[:, 0])
        return - 0.5 * ((torch.norm(z, dim=1) - 2) / 0.2) ** 2 - 0.5 * ((a - 2) / 0.3) ** 2 \
            + torch.log(1 + torch.exp(-4 * a / 0.09))


class NealsFunnel(Target):
    """Neal's funnel in two dimensions (target.py:65-104): v ~ N(0, 1 + v1shift), x | v ~ N(0, exp(v / 2) + v2shift).
    The reference evaluates it on the host (`.cpu()` on every term) and so returns a CPU tensor; this returns the same
    values on z's device."""

    def __init__(self, prop_scale=torch.tensor(20.), prop_shift=torch.tensor(-10.), v1shift=0., v2shift=0.):
        super().__init__(prop_scale, prop_shift)
        self.n_dims = 2
        self.max_log_prob = 0.
        self.v1shift, self.v2shift = v1shift, v2shift

    def log_prob(self, z):
        v, x = z[:, 0], z[:, 1]
        half_log_2pi = 0.5 * np.log(2 * np.pi)
        sv = torch.ones_like(v) + self.v1shift
        sx = torch.exp(0.5 * v) + self.v2shift
        return (-0.5 * (v / sv) ** 2 - torch.log(sv) - half_log_2pi) + (-0.5 * (x / sx) ** 2 - torch.log(sx) - half_log_2pi)
