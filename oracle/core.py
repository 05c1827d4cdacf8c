"""Oracle containers: `NormalizingFlow` (normflow/core.py:9-197), `MultiscaleFlow`
(:271-399) and variational dequantisation (normflow/nets/varquant.py:9-83).

Reference quirks kept: `forward_kld` accumulates log_q in fp32
(A.3) and the dequantisers mutate the accumulator they are handed, which makes the
per-column log-dets add up triangularly (A.2).  One documented deviation, shared with
the product: `categoricals=None` means "no dequantisation" instead of AttributeError (A.1).
"""
import math

import numpy as np
import torch
from torch import nn
import torch.nn.functional as F


class Dequantization(nn.Module):
    """varquant.py:9-51."""

    def __init__(self, alpha=1e-5, quants=256):
        super().__init__()
        self.alpha, self.quants = alpha, quants

    def forward(self, z, ldj, reverse=False):
        if not reverse:
            z, ldj = self.dequant(z, ldj)
            return self.sigmoid(z, ldj, reverse=True)
        z, ldj = self.sigmoid(z, ldj, reverse=False)
        z = z * self.quants
        ldj += np.log(self.quants) * np.prod(z.shape[1:])
        return torch.floor(z).clamp(min=0, max=self.quants - 1).to(torch.int32), ldj

    def sigmoid(self, z, ldj, reverse=False):
        dims = list(range(1, z.dim()))
        if not reverse:
            ldj += (-z - 2 * F.softplus(-z)).sum(dim=dims)
            return torch.sigmoid(z), ldj
        z = z * (1 - self.alpha) + 0.5 * self.alpha
        ldj += np.log(1 - self.alpha) * np.prod(z.shape[1:])
        ldj += (-torch.log(z) - torch.log(1 - z)).sum(dim=dims)
        return torch.log(z) - torch.log(1 - z), ldj

    def noise(self, z):
        return torch.rand_like(z).detach()

    def dequant(self, z, ldj):
        z = z.to(torch.float32)
        z = (z + self.noise(z)) / self.quants
        ldj -= np.log(self.quants) * np.prod(z.shape[1:])
        return z, ldj


class VariationalDequantization(Dequantization):
    """varquant.py:53-83: the uniform noise goes logit -> flows -> sigmoid first (fp64 noise)."""

    def __init__(self, var_flows, alpha=1e-5, quants=8):
        super().__init__(alpha=alpha, quants=quants)
        self.flows = nn.ModuleList(var_flows)

    def noise(self, z):
        return torch.rand_like(z, dtype=torch.double).detach()

    def dequant(self, z, ldj):
        z = z.to(torch.float32)
        u, ldj = self.sigmoid(self.noise(z), ldj, reverse=True)
        for f in self.flows:
            u, ld = f(u)
            ldj += ld
        u, ldj = self.sigmoid(u, ldj, reverse=False)
        z = (z + u) / self.quants
        ldj -= np.log(self.quants) * np.prod(z.shape[1:])
        return z, ldj


class NormalizingFlow(nn.Module):
    def __init__(self, q0, flows, p=None, categoricals=None, catlevels=None, catvdeqs=None):
        super().__init__()
        self.q0, self.flows, self.p = q0, nn.ModuleList(flows), p
        self.categoricals = categoricals
        if categoricals is not None:
            self.vdeqs, self.catlevels = nn.ModuleList(catvdeqs), catlevels

    def forward_kld(self, x, extended=False):                        # core.py:33-66
        log_qt = torch.zeros(len(x), device=x.device)
        log_q = torch.zeros(len(x), device=x.device)
        z = x
        if self.categoricals is not None:
            zc = x[:, self.categoricals]
            for k in range(len(self.categoricals)):
                zk, ld = self.vdeqs[k].forward(z=zc[:, k].view([zc.shape[0], 1]), ldj=log_q, reverse=False)
                zc[:, k] = zk.squeeze()
                log_qt += ld
            z[:, self.categoricals] = zc.float()
        log_q = log_qt
        for f in reversed(self.flows):
            z, ld = f.inverse(z)
            log_q += ld
        log_q += self.q0.log_prob(z)
        return -torch.mean(log_q)

    def reverse_kld(self, num_samples=1, beta=1.0, score_fn=True):   # core.py:68-101
        z, log_q_ = self.q0(num_samples)
        log_q = torch.zeros_like(log_q_) + log_q_
        for f in self.flows:
            z, ld = f(z)
            log_q = log_q - ld
        if not score_fn:
            z_ = z
            log_q = torch.zeros(len(z_), device=z_.device)
            for p in self.parameters():
                p.requires_grad_(False)
            for f in reversed(self.flows):
                z_, ld = f.inverse(z_)
                log_q = log_q + ld
            log_q = log_q + self.q0.log_prob(z_)
            for p in self.parameters():
                p.requires_grad_(True)
        return torch.mean(log_q) - beta * torch.mean(self.p.log_prob(z))

    def reverse_alpha_div(self, num_samples=1, alpha=1, dreg=False):   # core.py:102-142
        z, log_q = self.q0(num_samples)
        for f in self.flows:
            z, ld = f(z)
            log_q = log_q - ld
        log_p = self.p.log_prob(z)
        if not dreg:
            return np.sign(alpha - 1) * torch.logsumexp(alpha * (log_p - log_q), 0)
        w_const = torch.exp(log_p - log_q).detach()            # doubly reparametrised estimator (:121-136)
        z_ = z
        log_q = torch.zeros(len(z_), device=z_.device)
        for p in self.parameters():
            p.requires_grad_(False)
        for f in reversed(self.flows):
            z_, ld = f.inverse(z_)
            log_q = log_q + ld
        log_q = log_q + self.q0.log_prob(z_)
        for p in self.parameters():
            p.requires_grad_(True)
        w = torch.exp(log_p - log_q)
        w_alpha = w_const ** alpha
        w_alpha = w_alpha / torch.mean(w_alpha)
        weights = (1 - alpha) * w_alpha + alpha * w_alpha ** 2
        return -alpha * torch.mean(weights * torch.log(w))

    def sample(self, num_samples=1):                                 # core.py:144-168
        z, log_q = self.q0(num_samples)
        for f in self.flows:
            z, ld = f(z)
            log_q = log_q - ld
        if self.categoricals is not None:
            zc = z[:, self.categoricals]
            for k in range(len(self.categoricals)):
                zk, ld = self.vdeqs[k].forward(z=zc[:, k].view([zc.shape[0], 1]), ldj=log_q, reverse=True)
                zc[:, k] = zk.squeeze()
                log_q += ld
            z[:, self.categoricals] = zc.to(z.dtype)
        return z, log_q

    def log_prob(self, x):                                           # core.py:170-183
        log_q = torch.zeros(len(x), dtype=x.dtype, device=x.device)
        z = x
        for f in reversed(self.flows):
            z, ld = f.inverse(z)
            log_q = log_q + ld
        return log_q + self.q0.log_prob(z)


class MultiscaleFlow(nn.Module):
    def __init__(self, q0, flows, merges, transform=None, class_cond=True):
        super().__init__()
        self.q0 = nn.ModuleList(q0)
        self.flows = nn.ModuleList([nn.ModuleList(f) for f in flows])
        self.merges = nn.ModuleList(merges)
        self.transform, self.class_cond = transform, class_cond

    def forward_kld(self, x, y=None):
        return -torch.mean(self.log_prob(x, y))

    def log_prob(self, x, y=None):                                   # core.py:342-367
        log_q, z = 0, x
        if self.transform is not None:
            z, ld = self.transform.inverse(z)
            log_q = log_q + ld
        for i in range(len(self.q0) - 1, -1, -1):
            for f in reversed(self.flows[i]):
                z, ld = f.inverse(z)
                log_q = log_q + ld
            if i > 0:
                [z, z_], ld = self.merges[i - 1].inverse(z)
                log_q = log_q + ld
            else:
                z_ = z
            log_q = log_q + (self.q0[i].log_prob(z_, y) if self.class_cond else self.q0[i].log_prob(z_))
        return log_q

    def sample(self, num_samples=1, y=None):                         # core.py:308-340
        for i in range(len(self.q0)):
            z_, lq_ = self.q0[i](num_samples, y) if self.class_cond else self.q0[i](num_samples)
            if i == 0:
                log_q, z = lq_, z_
            else:
                log_q = log_q + lq_
                z, ld = self.merges[i - 1]([z, z_])
                log_q = log_q - ld
            for f in self.flows[i]:
                z, ld = f(z)
                log_q = log_q - ld
        if self.transform is not None:
            z, ld = self.transform(z)
            log_q = log_q - ld
        return z, log_q
