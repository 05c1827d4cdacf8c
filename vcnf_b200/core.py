"""Model containers with the reference's public API (normflow/core.py): the Python layer loops
stay on the host, every layer call lands in the sm_100a kernels.

Reference behaviours kept on purpose: `forward_kld` accumulates log q in fp32
whatever the model dtype (A.3), writes the dequantised columns back into the caller's `x`
(core.py:54) and hands the running accumulator to the dequantisers, which mutate it in place
(triangular accumulation, A.2).  Single intentional deviation: `categoricals=None` means "no
dequantisation" (the reference raises AttributeError at core.py:42,157, hazard A.1)."""
import numpy as np
import torch
import torch.nn as nn

from . import ops, utils
from .flows import mixing


class NormalizingFlow(nn.Module):
    def __init__(self, q0, flows, p=None, categoricals=None, catlevels=None, catvdeqs=None):
        super().__init__()
        self.q0 = q0
        self.flows = nn.ModuleList(flows)
        self.p = p
        self.categoricals = categoricals
        if categoricals is not None:
            self.vdeqs = nn.ModuleList(catvdeqs)
            self.catlevels = catlevels

    vdeq_noise_fn = None      # optional callable(B, n_columns, device) -> fp64 uniform noise (parity tests inject it)

    def _vdeq_plan(self, device):
        """(table [n, L, 2] fp64, quants [n] fp64, alpha) when every categorical column's dequantiser reduces to
        elementwise affine layers (nets.VariationalDequantization.affine_table), else None -> per-column route."""
        from .nets.varquant import VariationalDequantization
        vd = list(self.vdeqs)[:len(self.categoricals)]
        if not vd or any(type(v) is not VariationalDequantization for v in vd):
            return None
        if len({float(v.alpha) for v in vd}) != 1 or len({len(v.flows) for v in vd}) != 1:
            return None
        tabs, seen = [], {}
        for v in vd:
            key = tuple(id(f) for f in v.flows)        # dequantisers that share their layer objects share the table
            if key not in seen:
                t = v.affine_table()
                if t is None:
                    return None
                seen[key] = t.to(device=device, dtype=torch.float64)
            tabs.append(seen[key])
        tab = tabs[0].unsqueeze(0).expand(len(vd), -1, -1) if len(seen) == 1 else torch.stack(tabs)
        return tab.contiguous(), self._vdeq_quants_on(device), float(vd[0].alpha)

    def _vdeq_quants_on(self, device):
        """Number of levels per categorical column as a device fp64 vector (built once per device)."""
        vd = list(self.vdeqs)[:len(self.categoricals)]
        cache = self.__dict__.setdefault("_vdeq_quants", {})
        qkey = (str(device), tuple(float(v.quants) for v in vd))
        if qkey not in cache:
            cache[qkey] = torch.tensor(qkey[1], dtype=torch.float64, device=device)
        return cache[qkey]

    def _cat_index(self, device):
        """The categorical column list as a device index tensor, built once per device: indexing with the Python
        list uploads it from pageable host memory on every call (and cannot be captured into a CUDA graph)."""
        cache = self.__dict__.setdefault("_cat_idx", {})
        key = (str(device), tuple(self.categoricals))
        if key not in cache:
            cache[key] = torch.tensor(list(self.categoricals), dtype=torch.long, device=device)
        return cache[key]

    def _lu_prepare(self, absorb=False):
        """Density direction: the weights of all LULinearPermute layers of the model in one launch
        (flows/mixing.py precompute_lu_weights) -- they do not depend on the activations.  Returns (layers served,
        the 0-dim sum of their log-dets when `absorb`, else None)."""
        lus = self.__dict__.get("_lu_cache")
        if lus is None or lus[0] != len(self.flows):
            lus = (len(self.flows), [f.linear for f in self.flows if isinstance(f, mixing.LULinearPermute)])
            self.__dict__["_lu_cache"] = lus
        if len(lus[1]) < 2:
            return [], None
        return mixing.precompute_lu_weights(lus[1], absorb=absorb)

    def forward_kld(self, x, extended=False):
        """core.py:33-66."""
        log_qt = torch.zeros(len(x), device=x.device)
        log_q = torch.zeros(len(x), device=x.device)
        z = x
        zarr, zparr = [], []
        if self.categoricals is not None and len(self.categoricals) > 0:
            idx = self._cat_index(x.device)
            zc = x[:, idx]
            fused = self._vdeq_plan(x.device) if not extended else None
            if fused is not None:
                # every column through its dequantiser in ONE kernel (same triangular accumulation, fp32)
                tab, quants, alpha = fused
                n = len(self.categoricals)
                noise = (self.vdeq_noise_fn(len(x), n, x.device) if self.vdeq_noise_fn is not None
                         else torch.rand(len(x), n, dtype=torch.float64, device=x.device))
                zc, log_qt = ops.vdeq_fused(zc, noise, tab, quants, alpha)
            else:
                for vcn in range(len(self.categoricals)):
                    zct, log_det = self.vdeqs[vcn].forward(z=zc[:, vcn].view([zc.shape[0], 1]), ldj=log_q,
                                                           reverse=False)
                    zc[:, vcn] = zct.squeeze()
                    if extended:
                        zarr.append(z.cpu().detach().numpy())
                        zparr.append(log_det.cpu().detach().numpy())
                    log_qt += log_det
            z[:, idx] = zc.float()
        log_q = log_qt
        lus, ld_lu = self._lu_prepare(absorb=not extended)
        try:
            for i in range(len(self.flows) - 1, -1, -1):
                z, log_det = self.flows[i].inverse(z)
                if extended:
                    zarr.append(z.cpu().detach().numpy())
                    zparr.append(log_det.cpu().detach().numpy())
                if log_det is not mixing.ABSORBED:
                    log_q += log_det
        finally:
            mixing.clear_lu_weights(lus)
        if ld_lu is not None:
            log_q += ops.broadcast_scalar(ld_lu.to(torch.promote_types(ld_lu.dtype, z.dtype)), len(x))
        log_q += self.q0.log_prob(z)
        if extended:
            return -torch.mean(log_q), zarr, zparr, log_q
        return -torch.mean(log_q)

    def reverse_kld(self, num_samples=1, beta=1., score_fn=True, extended=False):
        """core.py:68-101."""
        z, log_q_ = self.q0(num_samples)
        log_q = torch.zeros_like(log_q_)
        log_q += log_q_
        zarr, zparr = [], []
        for flow in self.flows:
            z, log_det = flow(z)
            if extended:
                zarr.append(z.cpu().detach().numpy())
                zparr.append(log_det.cpu().detach().numpy())
            log_q -= log_det
        if not score_fn:
            z_ = z
            log_q = torch.zeros(len(z_), device=z_.device)
            utils.set_requires_grad(self, False)
            for i in range(len(self.flows) - 1, -1, -1):
                z_, log_det = self.flows[i].inverse(z_)
                log_q += log_det
            log_q += self.q0.log_prob(z_)
            utils.set_requires_grad(self, True)
        log_p = self.p.log_prob(z)
        if extended:
            return torch.mean(log_q) - beta * torch.mean(log_p), zarr, zparr, log_p, log_q
        return torch.mean(log_q) - beta * torch.mean(log_p)

    def reverse_alpha_div(self, num_samples=1, alpha=1, dreg=False, extended=False):
        """core.py:102-142."""
        z, log_q = self.q0(num_samples)
        zarr, zparr = [], []
        for flow in self.flows:
            z, log_det = flow(z)
            if extended:
                zarr.append(z.cpu().detach().numpy())
                zparr.append(log_det.cpu().detach().numpy())
            log_q -= log_det
        log_p = self.p.log_prob(z)
        if dreg:
            w_const = torch.exp(log_p - log_q).detach()
            z_ = z
            log_q = torch.zeros(len(z_), device=z_.device)
            utils.set_requires_grad(self, False)
            for i in range(len(self.flows) - 1, -1, -1):
                z_, log_det = self.flows[i].inverse(z_)
                log_q += log_det
            log_q += self.q0.log_prob(z_)
            utils.set_requires_grad(self, True)
            w = torch.exp(log_p - log_q)
            w_alpha = w_const ** alpha
            w_alpha = w_alpha / torch.mean(w_alpha)
            weights = (1 - alpha) * w_alpha + alpha * w_alpha ** 2
            loss = -alpha * torch.mean(weights * torch.log(w))
        else:
            loss = np.sign(alpha - 1) * torch.logsumexp(alpha * (log_p - log_q), 0)
        if extended:
            return loss, zarr, zparr
        return loss

    def sample(self, num_samples=1):
        """core.py:144-168."""
        z, log_q = self.q0(num_samples)
        lus, ld_lu = [], None
        if not torch.is_grad_enabled():
            # every LU layer is inverted once per pass and none depends on the samples: one launch for all of them, their
            # log-dets as one model-level term (flows/mixing.py precompute_lu_inverses)
            cache = self.__dict__.get("_lu_cache")
            if cache is None or cache[0] != len(self.flows):
                cache = (len(self.flows), [f.linear for f in self.flows if isinstance(f, mixing.LULinearPermute)])
                self.__dict__["_lu_cache"] = cache
            if len(cache[1]) >= 2:
                lus, ld_lu = mixing.precompute_lu_inverses(cache[1], absorb=True)
        try:
            for flow in self.flows:
                z, log_det = flow(z)
                if log_det is not mixing.ABSORBED:
                    log_q -= log_det
        finally:
            mixing.clear_lu_weights(lus)
        if ld_lu is not None:       # sampling direction: log|det| of each layer is -log|det L U|, and log_q -= log_det
            log_q += ops.broadcast_scalar(ld_lu.to(torch.promote_types(ld_lu.dtype, z.dtype)), len(z))
        if self.categoricals is not None and len(self.categoricals) > 0:
            idx = self._cat_index(z.device)
            zc = z[:, idx]
            plain = all(type(v).__name__ in ("VariationalDequantization", "Dequantization")
                        for v in list(self.vdeqs)[:len(self.categoricals)])
            if plain and not torch.is_grad_enabled() and log_q.dtype == zc.dtype and log_q.is_contiguous():
                # floor / clamp to the level index for all columns in one kernel; log_q updated in place with the
                # reference's aliasing (`log_q += log_det` where log_det IS log_q, core.py:160-164)
                quants = self._vdeq_quants_on(z.device)
                zc = ops.vdeq_decode(zc, quants, log_q)
            else:
                for vct in range(len(self.categoricals)):
                    zct, log_det = self.vdeqs[vct].forward(z=zc[:, vct].view([zc.shape[0], 1]), ldj=log_q,
                                                           reverse=True)
                    zc[:, vct] = zct.squeeze()
                    log_q += log_det
            z[:, idx] = zc.to(z.dtype)
        return z, log_q

    def log_prob(self, x):
        """core.py:170-183."""
        log_q = torch.zeros(len(x), dtype=x.dtype, device=x.device)
        z = x
        lus, ld_lu = self._lu_prepare(absorb=True)
        try:
            for i in range(len(self.flows) - 1, -1, -1):
                z, log_det = self.flows[i].inverse(z)
                if log_det is not mixing.ABSORBED:
                    log_q += log_det
        finally:
            mixing.clear_lu_weights(lus)
        if ld_lu is not None:
            log_q += ops.broadcast_scalar(ld_lu.to(torch.promote_types(ld_lu.dtype, z.dtype)), len(x))
        log_q += self.q0.log_prob(z)
        return log_q

    def save(self, path):
        torch.save(self.state_dict(), path)

    def load(self, path):
        self.load_state_dict(torch.load(path))


class ClassCondFlow(nn.Module):
    """Flow whose base distribution is conditioned on a class label (core.py:200-268): the layer loops of
    NormalizingFlow with `q0(num_samples, y)` / `q0.log_prob(z, y)` at the latent end (ClassCondDiagGaussian:
    distributions/base.py, one kernel per direction)."""

    def __init__(self, q0, flows):
        super().__init__()
        self.q0 = q0
        self.flows = nn.ModuleList(flows)

    def log_prob(self, x, y):
        log_q = torch.zeros(len(x), dtype=x.dtype, device=x.device)
        z = x
        for flow in reversed(self.flows):
            z, log_det = flow.inverse(z)
            log_q += log_det
        log_q += self.q0.log_prob(z, y)
        return log_q

    def forward_kld(self, x, y):
        return -torch.mean(self.log_prob(x, y))

    def sample(self, num_samples=1, y=None):
        z, log_q = self.q0(num_samples, y)
        for flow in self.flows:
            z, log_det = flow(z)
            log_q -= log_det
        return z, log_q

    def save(self, path):
        torch.save(self.state_dict(), path)

    def load(self, path):
        self.load_state_dict(torch.load(path))


class MultiscaleFlow(nn.Module):
    """Multiscale (RealNVP / Glow) container, core.py:271-399."""

    def __init__(self, q0, flows, merges, transform=None, class_cond=True):
        super().__init__()
        self.q0 = nn.ModuleList(q0)
        self.num_levels = len(self.q0)
        self.flows = torch.nn.ModuleList([nn.ModuleList(flow) for flow in flows])
        self.merges = torch.nn.ModuleList(merges)
        self.transform = transform
        self.class_cond = class_cond

    def forward_kld(self, x, y=None):
        return -torch.mean(self.log_prob(x, y))

    def forward(self, x, y=None):
        return -self.log_prob(x, y)

    def sample(self, num_samples=1, y=None, temperature=None):
        if temperature is not None:
            self.set_temperature(temperature)
        for i in range(len(self.q0)):
            if self.class_cond:
                z_, log_q_ = self.q0[i](num_samples, y)
            else:
                z_, log_q_ = self.q0[i](num_samples)
            if i == 0:
                log_q, z = log_q_, z_
            else:
                log_q += log_q_
                z, log_det = self.merges[i - 1]([z, z_])
                log_q -= log_det
            for flow in self.flows[i]:
                z, log_det = flow(z)
                log_q -= log_det
        if self.transform is not None:
            z, log_det = self.transform(z)
            log_q -= log_det
        if temperature is not None:
            self.reset_temperature()
        return z, log_q

    def log_prob(self, x, y):
        log_q = 0
        z = x
        if self.transform is not None:
            z, log_det = self.transform.inverse(z)
            log_q += log_det
        for i in range(len(self.q0) - 1, -1, -1):
            for j in range(len(self.flows[i]) - 1, -1, -1):
                z, log_det = self.flows[i][j].inverse(z)
                log_q += log_det
            if i > 0:
                [z, z_], log_det = self.merges[i - 1].inverse(z)
                log_q += log_det
            else:
                z_ = z
            if self.class_cond:
                log_q += self.q0[i].log_prob(z_, y)
            else:
                log_q += self.q0[i].log_prob(z_)
        return log_q

    def save(self, path):
        torch.save(self.state_dict(), path)

    def load(self, path):
        self.load_state_dict(torch.load(path))

    def set_temperature(self, temperature):
        for q0 in self.q0:
            if hasattr(q0, "temperature"):
                q0.temperature = temperature
            else:
                raise NotImplementedError("One base function does not support temperature annealed sampling")

    def reset_temperature(self):
        self.set_temperature(None)
