"""(Variational) dequantisation of categorical columns (reference: normflow/nets/varquant.py:9-83).
The accumulator `ldj` is mutated in place exactly like the reference (this is what makes the
per-column log-dets add up triangularly in NormalizingFlow.forward_kld).  The
flow stack inside runs on the bijector kernels; the logit / sigmoid glue on [B, 1] columns is a
few library elementwise ops (a fused per-column kernel is a later row)."""
import numpy as np
import torch
from torch import nn
import torch.nn.functional as F


class Dequantization(nn.Module):
    def __init__(self, alpha=1e-5, quants=256):
        super().__init__()
        self.alpha = alpha
        self.quants = quants

    def forward(self, z, ldj, reverse=False):
        if not reverse:
            z, ldj = self.dequant(z, ldj)
            z, ldj = self.sigmoid(z, ldj, reverse=True)
        else:
            z, ldj = self.sigmoid(z, ldj, reverse=False)
            z = z * self.quants
            ldj += np.log(self.quants) * np.prod(z.shape[1:])
            z = torch.floor(z).clamp(min=0, max=self.quants - 1).to(torch.int32)
        return z, ldj

    def sigmoid(self, z, ldj, reverse=False):
        dims = list(range(1, len(z.shape)))
        if not reverse:
            ldj += (-z - 2 * F.softplus(-z)).sum(dim=dims)
            z = torch.sigmoid(z)
        else:
            z = z * (1 - self.alpha) + 0.5 * self.alpha
            ldj += np.log(1 - self.alpha) * np.prod(z.shape[1:])
            ldj += (-torch.log(z) - torch.log(1 - z)).sum(dim=dims)
            z = torch.log(z) - torch.log(1 - z)
        return z, ldj

    def dequant(self, z, ldj):
        z = z.to(torch.float32)
        z = z + torch.rand_like(z).detach()
        z = z / self.quants
        ldj -= np.log(self.quants) * np.prod(z.shape[1:])
        return z, ldj


class VariationalDequantization(Dequantization):
    def __init__(self, var_flows, alpha=1e-5, quants=8):
        super().__init__(alpha=alpha, quants=quants)
        self.flows = nn.ModuleList(var_flows)

    def dequant(self, z, ldj):
        z = z.to(torch.float32)
        deq_noise = torch.rand_like(z, dtype=torch.double).detach()       # varquant.py:70
        deq_noise, ldj = self.sigmoid(deq_noise, ldj, reverse=True)
        for flow in self.flows:
            deq_noise, log_det = flow(deq_noise)
            ldj += log_det
        deq_noise, ldj = self.sigmoid(deq_noise, ldj, reverse=False)
        z = (z + deq_noise) / self.quants
        ldj -= np.log(self.quants) * np.prod(z.shape[1:])
        return z, ldj
