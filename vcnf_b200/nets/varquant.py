"""(Variational) dequantisation of categorical columns -- public classes of normflow/nets/varquant.py:9-83.

Hot path: `NormalizingFlow.forward_kld / sample` do not call these modules column by column; they ask
`affine_table()` for each dequantiser's per-layer (log-scale, shift) pairs and run ALL columns in one kernel
(vcnf_vdeq_fwd / _bwd / _decode, csrc/vdeq_kernels.cu).  The methods below are the generic route for stacks the
kernel does not cover (layers that are not elementwise-affine on a [B, 1] column, or an ActNorm still waiting for
its data-dependent initialisation): same arithmetic as the reference, written with the log-det bookkeeping in two
helpers; the accumulator `ldj` is updated IN PLACE because core.py:47-52 depends on that aliasing (hazard A.2)."""
import math

import torch
from torch import nn
import torch.nn.functional as F

# noise source of the generic route (tests substitute a host-seeded draw to compare with the CPU reference)
_rand_like = torch.rand_like


def _row_sum(t):
    return t.flatten(1).sum(dim=1)


def _logit_(z, ldj, alpha):
    """z -> logit(z (1 - alpha) + alpha / 2); adds the log-det to ldj in place (varquant.py:38-42)."""
    z = z * (1 - alpha) + 0.5 * alpha
    n_elem = math.prod(z.shape[1:])
    ldj += math.log(1 - alpha) * n_elem
    lz, l1z = torch.log(z), torch.log(1 - z)
    ldj += _row_sum(-lz - l1z)
    return lz - l1z, ldj


def _sigmoid_(z, ldj):
    """z -> sigmoid(z); adds the log-det -z - 2 softplus(-z) to ldj in place (varquant.py:35-37)."""
    ldj += _row_sum(-z - 2 * F.softplus(-z))
    return torch.sigmoid(z), ldj


class Dequantization(nn.Module):
    def __init__(self, alpha=1e-5, quants=256):
        super().__init__()
        self.alpha = alpha
        self.quants = quants

    def sigmoid(self, z, ldj, reverse=False):
        return _logit_(z, ldj, self.alpha) if reverse else _sigmoid_(z, ldj)

    def dequant(self, z, ldj):
        z = z.to(torch.float32)
        z = (z + _rand_like(z).detach()) / self.quants
        ldj -= math.log(self.quants) * math.prod(z.shape[1:])
        return z, ldj

    def forward(self, z, ldj, reverse=False):
        if reverse:                                    # sampling: continuous -> level index (varquant.py:26-30)
            z, ldj = _sigmoid_(z, ldj)
            z = z * self.quants
            ldj += math.log(self.quants) * math.prod(z.shape[1:])
            return torch.floor(z).clamp(min=0, max=self.quants - 1).to(torch.int32), ldj
        z, ldj = self.dequant(z, ldj)
        return _logit_(z, ldj, self.alpha)


class VariationalDequantization(Dequantization):
    def __init__(self, var_flows, alpha=1e-5, quants=8):
        super().__init__(alpha=alpha, quants=quants)
        self.flows = nn.ModuleList(var_flows)

    def dequant(self, z, ldj):
        z = z.to(torch.float32)
        u = _rand_like(z, dtype=torch.double).detach()             # fp64 noise whatever the model (varquant.py:70)
        u, ldj = _logit_(u, ldj, self.alpha)
        for flow in self.flows:
            u, log_det = flow(u)
            ldj += log_det
        u, ldj = _sigmoid_(u, ldj)
        z = (z + u) / self.quants
        ldj -= math.log(self.quants) * math.prod(z.shape[1:])
        return z, ldj

    def affine_table(self):
        """[L, 2] tensor of (log-scale, shift) per layer when every layer acts on a [B, 1] column as
        v -> v exp(a) + c with log-det a, else None.  Differentiable w.r.t. the layers' parameters."""
        from ..flows.affine.coupling import AffineConstFlow, MaskedAffineFlow
        from ..flows.normalization import ActNorm
        a, c = [], []
        for flow in self.flows:
            if isinstance(flow, AffineConstFlow):
                if flow.s.numel() != 1 or flow.t.numel() != 1:
                    return None
                if isinstance(flow, ActNorm) and not flow.initialised():
                    return None                       # first batch: statistics layer by layer (normalization.py:22-27)
                a.append(flow.s.reshape(1))
                c.append(flow.t.reshape(1))
            elif isinstance(flow, MaskedAffineFlow):
                if flow.b_cpu.numel() != 1 or float(flow.b_cpu) not in (0.0, 1.0):
                    return None
                ref = next(flow.parameters(), flow.b)
                z0 = torch.zeros(1, 1, dtype=ref.dtype, device=ref.device)     # b z == 0: constant conditioner input
                sv = flow.s(z0).reshape(1) if flow._has_s else z0.reshape(1)
                tv = flow.t(z0).reshape(1) if flow._has_t else z0.reshape(1)
                if float(flow.b_cpu) == 1.0:
                    # pass-through, b z + 0 * (...) (coupling.py:209): the conditioners stay in the graph with a
                    # ZERO gradient like the reference's (an optimiser with weight decay treats None differently)
                    a.append(sv * 0)
                    c.append(tv * 0)
                    continue
                nan = torch.full_like(sv, float("nan"))
                a.append(torch.where(torch.isfinite(sv), sv, nan))              # coupling.py:205-208
                c.append(torch.where(torch.isfinite(tv), tv, nan))
            else:
                return None
        if not a:
            return torch.zeros(0, 2, dtype=torch.float64)
        return torch.stack([torch.cat(a), torch.cat(c)], dim=1)
