"""Mixing layers (reference: normflow/flows/mixing.py): Permute, Invertible1x1Conv,
LULinearPermute.  The per-sample / per-pixel D x D maps run in `vcnf_small_linear`."""
import numpy as np
import torch
from torch import nn
from torch.nn import functional as F, init

from .base import Flow
from .. import ops


_I32_CACHE = {}


def _i32(t):
    """int32 copy of an index buffer for the kernels, converted once per (storage, version, device) instead of
    once per call (one tiny launch per layer and direction otherwise)."""
    key = (t.data_ptr(), t._version, t.device, tuple(t.shape))
    hit = _I32_CACHE.get(id(t))
    if hit is not None and hit[0] == key:
        return hit[1]
    if len(_I32_CACHE) > 4096:
        _I32_CACHE.clear()
    out = t.to(torch.int32).contiguous()
    _I32_CACHE[id(t)] = (key, out, t)          # holding `t` keeps its id from being reused
    return out


class Permute(Flow):
    """Channel permutation (mixing.py:10-54).  'swap': forward splits at C//2, inverse at
    (C+1)//2 (hazard A.13); 'shuffle': fixed random permutation (buffers perm, inv_perm)."""

    def __init__(self, num_channels, mode="shuffle"):
        super().__init__()
        self.mode = mode
        self.num_channels = num_channels
        if self.mode == "shuffle":
            perm = torch.randperm(self.num_channels)
            inv_perm = torch.empty_like(perm).scatter_(dim=0, index=perm, src=torch.arange(self.num_channels))
            self.register_buffer("perm", perm)
            self.register_buffer("inv_perm", inv_perm)
        self._idx_cache = {}

    def _swap_index(self, cut, device):
        key = (cut, str(device))
        if key not in self._idx_cache:
            c = self.num_channels
            fwd = torch.cat([torch.arange(cut, c), torch.arange(0, cut)]).to(torch.int32).to(device)
            self._idx_cache[key] = (fwd, torch.argsort(fwd.long()).to(torch.int32))
        return self._idx_cache[key]

    def _apply_perm(self, z, inverse):
        if self.mode == "shuffle":
            p, q = (self.inv_perm, self.perm) if inverse else (self.perm, self.inv_perm)
            return ops.permute_channels(z, _i32(p), _i32(q))
        if self.mode == "swap":
            cut = (self.num_channels + 1) // 2 if inverse else self.num_channels // 2
            p, q = self._swap_index(cut, z.device)
            return ops.permute_channels(z, p, q)
        raise NotImplementedError("The mode " + self.mode + " is not implemented.")

    def forward(self, z):
        return self._apply_perm(z, False), 0

    def inverse(self, z):
        return self._apply_perm(z, True), 0


class Invertible1x1Conv(Flow):
    """Glow's invertible 1x1 convolution (mixing.py:57-128).  W (or its inverse, with the
    triangular factors inverted in fp64 as the reference does, :91-96) is assembled from the C x C
    parameters by a handful of tiny library ops; the per-pixel C x C matvec over the NCHW tensor
    is the `vcnf_small_linear` kernel."""

    def __init__(self, num_channels, use_lu=False):
        super().__init__()
        self.num_channels = num_channels
        self.use_lu = use_lu
        Q = torch.linalg.qr(torch.randn(self.num_channels, self.num_channels))[0]
        if use_lu:
            P, L, U = torch.linalg.lu(Q)
            self.register_buffer("P", P)
            self.L = nn.Parameter(L)
            S = U.diag()
            self.register_buffer("sign_S", torch.sign(S))
            self.log_S = nn.Parameter(torch.log(torch.abs(S)))
            self.U = nn.Parameter(torch.triu(U, diagonal=1))
            self.register_buffer("eye", torch.diag(torch.ones(self.num_channels)))
        else:
            self.W = nn.Parameter(Q)

    def _assemble_W(self, inverse=False):
        L = torch.tril(self.L, diagonal=-1) + self.eye
        U = torch.triu(self.U, diagonal=1) + torch.diag(self.sign_S * torch.exp(self.log_S))
        if inverse:
            if self.log_S.dtype == torch.float64:
                L_inv, U_inv = torch.inverse(L), torch.inverse(U)
            else:
                L_inv = torch.inverse(L.double()).type(self.log_S.dtype)
                U_inv = torch.inverse(U.double()).type(self.log_S.dtype)
            return U_inv @ L_inv @ self.P.t()
        return self.P @ L @ U

    def forward(self, z):
        if self.use_lu:
            W = self._assemble_W(inverse=True)
            log_det = -torch.sum(self.log_S)
        else:
            W_dtype = self.W.dtype
            W = torch.inverse(self.W) if W_dtype == torch.float64 else torch.inverse(self.W.double()).type(W_dtype)
            log_det = -torch.slogdet(self.W)[1]
        z_ = ops.small_linear(z, W)
        return z_, log_det * z.size(2) * z.size(3)

    def inverse(self, z):
        if self.use_lu and z.is_cuda and self.num_channels <= ops.GLOW_W_MAX and self.L.dtype == self.P.dtype:
            # density direction (the training path): W and H W sum(log_S) in one launch each way
            W, log_det = ops.glow_w(self.L, self.U, self.log_S, self.P, self.sign_S, z.size(2) * z.size(3))
            return ops.small_linear(z, W), log_det
        if self.use_lu:
            W = self._assemble_W()
            log_det = torch.sum(self.log_S)
        else:
            W = self.W
            log_det = torch.slogdet(self.W)[1]
        z_ = ops.small_linear(z, W)
        return z_, log_det * z.size(2) * z.size(3)


class _Permutation(Flow):
    """Fixed feature permutation (mixing.py:203-233); used through LULinearPermute."""

    def __init__(self, permutation, dim=1):
        if permutation.ndimension() != 1:
            raise ValueError("Permutation must be a 1D tensor.")
        super().__init__()
        self._dim = dim
        self.register_buffer("_permutation", permutation)

    @property
    def _inverse_permutation(self):
        return torch.argsort(self._permutation)

    def forward(self, inputs, context=None):
        out = ops.permute_channels(inputs, _i32(self._permutation), _i32(self._inverse_permutation))
        return out, torch.zeros(inputs.shape[0], dtype=inputs.dtype, device=inputs.device)

    def inverse(self, inputs, context=None):
        out = ops.permute_channels(inputs, _i32(self._inverse_permutation), _i32(self._permutation))
        return out, torch.zeros(inputs.shape[0], dtype=inputs.dtype, device=inputs.device)


class _RandomPermutation(_Permutation):
    def __init__(self, features, dim=1):
        super().__init__(torch.randperm(features), dim)


class _LULinear(Flow):
    """W = L U with L unit lower triangular and diag(U) = softplus(d) + eps (mixing.py:352-462).
    Parameters: bias, lower_entries, upper_entries, unconstrained_upper_diag."""

    def __init__(self, features, using_cache=False, identity_init=True, eps=1e-3):
        super().__init__()
        self.features = features
        self.eps = eps
        self.bias = nn.Parameter(torch.zeros(features))
        self.lower_indices = np.tril_indices(features, k=-1)
        self.upper_indices = np.triu_indices(features, k=1)
        self.diag_indices = np.diag_indices(features)
        n_tri = ((features - 1) * features) // 2
        self.lower_entries = nn.Parameter(torch.zeros(n_tri))
        self.upper_entries = nn.Parameter(torch.zeros(n_tri))
        self.unconstrained_upper_diag = nn.Parameter(torch.zeros(features))
        init.zeros_(self.bias)
        if identity_init:
            init.constant_(self.unconstrained_upper_diag, np.log(np.exp(1 - self.eps) - 1))
        else:
            stdv = 1.0 / np.sqrt(self.features)
            init.uniform_(self.lower_entries, -stdv, stdv)
            init.uniform_(self.upper_entries, -stdv, stdv)
            init.uniform_(self.unconstrained_upper_diag, -stdv, stdv)

    @property
    def upper_diag(self):
        return F.softplus(self.unconstrained_upper_diag) + self.eps

    def _index_tensors(self, device):
        """Triangular index sets as DEVICE tensors, built once per device: indexing with the numpy arrays
        would stage a host->device copy on every call (and is illegal inside CUDA-graph capture)."""
        cache = getattr(self, "_idx_cache", None)
        if cache is None or cache[0] != device:
            mk = lambda a: torch.as_tensor(np.asarray(a), dtype=torch.long, device=device)
            cache = (device, tuple(mk(a) for a in (self.lower_indices[0], self.lower_indices[1],
                                                    self.upper_indices[0], self.upper_indices[1],
                                                    self.diag_indices[0])))
            self._idx_cache = cache
        return cache[1]

    def _create_lower_upper(self):
        li, lj, ui, uj, di = self._index_tensors(self.lower_entries.device)
        lower = self.lower_entries.new_zeros(self.features, self.features)
        lower = lower.index_put((li, lj), self.lower_entries)
        lower = lower.index_put((di, di), self.lower_entries.new_ones(self.features))
        upper = self.upper_entries.new_zeros(self.features, self.features)
        upper = upper.index_put((ui, uj), self.upper_entries)
        upper = upper.index_put((di, di), self.upper_diag)
        return lower, upper

    def weight(self):
        lower, upper = self._create_lower_upper()
        return lower @ upper

    def weight_inverse(self):
        """U^-1 L^-1 via two triangular solves against the identity, in fp64 for fp32 models."""
        lower, upper = self._create_lower_upper()
        dt = lower.dtype
        eye = torch.eye(self.features, dtype=torch.float64, device=lower.device)
        li = torch.linalg.solve_triangular(lower.double(), eye, upper=False, unitriangular=True)
        wi = torch.linalg.solve_triangular(upper.double(), li, upper=True)
        return wi.to(dt)

    def logabsdet(self):
        return torch.sum(torch.log(self.upper_diag))

    def _fused(self):
        """Parameters on a CUDA device and D <= 64: W and log|det| come from one fused kernel."""
        return self.lower_entries.is_cuda and self.features <= 64

    def forward(self, inputs, context=None, in_idx=None):
        pre = self.__dict__.pop("_pre", None)                   # formed by precompute_lu_weights for this pass
        if pre is not None:
            weight, logabsdet = pre
        elif self._fused():
            weight, logabsdet = ops.lu_compose(self.lower_entries, self.upper_entries, self.unconstrained_upper_diag,
                                               self.eps)
        else:
            weight, logabsdet = self.weight(), self.logabsdet()
        out = ops.small_linear(inputs, weight, self.bias, in_idx=in_idx)
        if logabsdet is None:                                   # absorbed into the model-level sum (precompute_lu_weights)
            return out, None
        return out, ops.broadcast_scalar(logabsdet.to(torch.promote_types(logabsdet.dtype, inputs.dtype)), out.shape[0])

    def inverse(self, inputs, context=None, out_idx=None):
        needs_grad = torch.is_grad_enabled() and (inputs.requires_grad or self.lower_entries.requires_grad)
        pre = self.__dict__.pop("_pre_inv", None)               # formed by precompute_lu_inverses for this pass
        if pre is not None and not needs_grad and pre[0].dtype == inputs.dtype:
            out = ops.small_linear(inputs, pre[0], self.bias, out_idx=out_idx, sub_bias_first=True)
            if pre[1] is None:                                  # -log|det| absorbed into the model-level sum
                return out, None
            return out, ops.broadcast_scalar(-pre[1].to(torch.promote_types(pre[1].dtype, inputs.dtype)), out.shape[0])
        if self._fused() and not needs_grad:
            # sampling without a gradient: (L U)^-1 and log|det| from one kernel instead of ~25 tiny launches
            winv, logabsdet = ops.lu_inverse(self.lower_entries, self.upper_entries, self.unconstrained_upper_diag,
                                             self.eps)
            if winv.dtype != inputs.dtype:
                winv, logabsdet = winv.to(inputs.dtype), logabsdet.to(inputs.dtype)
        else:
            winv, logabsdet = self.weight_inverse(), self.logabsdet()
        out = ops.small_linear(inputs, winv, self.bias, out_idx=out_idx, sub_bias_first=True)
        return out, ops.broadcast_scalar(-logabsdet.to(torch.promote_types(logabsdet.dtype, inputs.dtype)), out.shape[0])


ABSORBED = 0.0      # log-det placeholder of a layer whose log|det W| the caller already holds (see precompute_lu_weights)


def precompute_lu_weights(layers, absorb=False):
    """W = L U and log|det W| of every `_LULinear` in `layers` with ONE launch per group of equal (D, dtype, eps)
    (ops.lu_compose_batched) instead of one single-CTA launch per layer; each layer's next density-direction call
    (`_LULinear.forward`) picks its pair up.  The adjoint is one launch as well, after the first layer's backward.
    Returns (layers served, total): with `absorb` the served layers return ABSORBED instead of a per-sample log-det
    vector and `total` is the 0-dim sum of their log|det W| (one broadcast add for the model instead of a copy, an add
    and two batch reductions in the backward per layer); otherwise total is None.  The caller clears leftovers with
    `clear_lu_weights`."""
    total = None
    groups = {}
    for m in layers:
        if m._fused() and m.features > 0:
            key = (m.features, m.lower_entries.dtype, float(m.eps), m.lower_entries.device)
            groups.setdefault(key, []).append(m)
    served = []
    for (_, _, eps, _), ms in groups.items():
        for k in range(0, len(ms), ops.LU_BATCH_MAX):
            chunk = ms[k:k + ops.LU_BATCH_MAX]
            if len(chunk) < 2:
                continue
            pairs = ops.lu_compose_batched([(m.lower_entries, m.upper_entries, m.unconstrained_upper_diag) for m in chunk],
                                           eps)
            for m, pr in zip(chunk, pairs):
                m.__dict__["_pre"] = (pr[0], None) if absorb else pr
            if absorb:
                part = torch.stack([pr[1] for pr in pairs]).sum()
                total = part if total is None else total + part
            served += chunk
    return served, total


def precompute_lu_inverses(layers, absorb=False):
    """Sampling direction without gradients: (L U)^-1 and log|det| of every `_LULinear` in `layers` with ONE launch per
    group of equal (D, dtype, eps) (ops.lu_inverse_batched) instead of one launch per layer; each layer's next
    `_LULinear.inverse` call picks its pair up.  Returns (layers served, total) like precompute_lu_weights: with `absorb`
    the served layers return ABSORBED and `total` is the 0-dim sum of their log|det W| (the sampling direction's log-det
    is its negative)."""
    total = None
    groups = {}
    for m in layers:
        if m._fused() and m.features > 0:
            key = (m.features, m.lower_entries.dtype, float(m.eps), m.lower_entries.device)
            groups.setdefault(key, []).append(m)
    served = []
    for (_, _, eps, _), ms in groups.items():
        for k in range(0, len(ms), ops.LU_BATCH_MAX):
            chunk = ms[k:k + ops.LU_BATCH_MAX]
            if len(chunk) < 2:
                continue
            pairs, lds = ops.lu_inverse_batched([(m.lower_entries, m.upper_entries, m.unconstrained_upper_diag)
                                                 for m in chunk], eps)
            for m, pr in zip(chunk, pairs):
                m.__dict__["_pre_inv"] = (pr[0], None) if absorb else pr
            if absorb:
                part = lds.sum()
                total = part if total is None else total + part
            served += chunk
    return served, total


def clear_lu_weights(layers):
    for m in layers:
        m.__dict__.pop("_pre", None)
        m.__dict__.pop("_pre_inv", None)


class LULinearPermute(Flow):
    """Permutation + LU-parametrised linear map (mixing.py:465-492), one kernel per direction:
    inverse (density): y = (L U) x[perm] + bias;  forward (sampling): x = ((L U)^-1 (y - bias))
    scattered back through the permutation.  The D x D product L U (and its inverse) is formed from
    the parameters by tiny library ops under autograd; the reference applies U then L (two
    F.linear calls, :398-410) and two triangular solves (:412-429) -- same map, rounding differs."""

    def __init__(self, num_channels, identity_init=True):
        super().__init__()
        self.permutation = _RandomPermutation(num_channels)
        self.linear = _LULinear(num_channels, identity_init=identity_init)

    def forward(self, z):
        perm = _i32(self.permutation._permutation)
        # out[:, perm[o]] = v[o]  ==  v[:, argsort(perm)]
        z, log_det = self.linear.inverse(z, out_idx=perm)
        if log_det is None:
            return z, ABSORBED
        return z, log_det.view(-1)

    def inverse(self, z):
        perm = _i32(self.permutation._permutation)
        z, log_det = self.linear(z, in_idx=perm)
        if log_det is None:
            return z, ABSORBED
        return z, log_det.view(-1)
