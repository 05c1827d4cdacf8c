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
