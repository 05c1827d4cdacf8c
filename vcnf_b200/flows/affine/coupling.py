"""Affine coupling family (reference: normflow/flows/affine/coupling.py) on sm_100a kernels."""
import numpy as np
import torch
from torch import nn

from ..base import Flow
from ..reshape import Split, Merge
from ... import ops


class AffineConstFlow(Flow):
    """z*exp(s)+t with learned per-feature constants (coupling.py:10-53); log_det is the 0-dim
    tensor prod(spatial)*sum(s).  Elementwise part + its parameter-gradient reduction run in
    `vcnf_const_affine_{fwd,bwd}`."""

    def __init__(self, shape, scale=True, shift=True):
        super().__init__()
        if scale:
            self.s = nn.Parameter(torch.zeros(shape)[None])
        else:
            self.register_buffer("s", torch.zeros(shape)[None])
        if shift:
            self.t = nn.Parameter(torch.zeros(shape)[None])
        else:
            self.register_buffer("t", torch.zeros(shape)[None])
        self.n_dim = self.s.dim()
        self.batch_dims = [i for i, n in enumerate(self.s.shape) if n == 1]

    def _spatial(self, z):
        if len(self.batch_dims) > 1:
            return int(np.prod([z.size(i) for i in self.batch_dims[1:]]))
        return 1

    def _apply_kernel(self, z, inverse):
        S = self._spatial(z)
        # broadcast dims must trail the parameter dims ((C,1,1) on NCHW, (D,) on [B,D])
        lead = [n for n in self.s.shape[1:] if n != 1]
        if list(self.s.shape[1:1 + len(lead)]) != lead or z.dim() != self.s.dim():
            raise NotImplementedError("AffineConstFlow: parameter shape %s not supported for input %s"
                                      % (tuple(self.s.shape), tuple(z.shape)))
        return ops.const_affine(z, self.s, self.t, S, inverse), S

    def forward(self, z):
        z_, S = self._apply_kernel(z, False)
        return z_, S * torch.sum(self.s)

    def inverse(self, z):
        z_, S = self._apply_kernel(z, True)
        return z_, -S * torch.sum(self.s)


class AffineCoupling(Flow):
    """RealNVP coupling on a pre-split pair [z1, z2] (coupling.py:94-168)."""

    def __init__(self, param_map, scale=True, scale_map="exp"):
        super().__init__()
        self.add_module("param_map", param_map)
        self.scale = scale
        self.scale_map = scale_map

    def _map_code(self):
        if not self.scale:
            return ops.SCALE_MAPS["none"]
        if self.scale_map not in ("exp", "sigmoid", "sigmoid_inv"):
            raise NotImplementedError("This scale map is not implemented.")
        return ops.SCALE_MAPS[self.scale_map]

    def _run(self, z, inverse):
        z1, z2 = z
        param = self.param_map(z1)
        code = self._map_code()
        out, ld = ops.affine_coupling(z2, param, 0, z2.shape[1], code, inverse)
        return [z1, out], (ld if self.scale else 0)

    def forward(self, z):
        return self._run(z, False)

    def inverse(self, z):
        return self._run(z, True)


class MaskedAffineFlow(Flow):
    """b*z + (1-b)*(z*exp(s(b*z)) + t(b*z)) (coupling.py:171-222), one fused kernel after the
    two conditioner nets (`vcnf_masked_affine_{fwd,bwd}`)."""

    def __init__(self, b, t=None, s=None):
        super().__init__()
        self.b_cpu = b.view(1, *b.size())
        self.register_buffer("b", self.b_cpu)
        self._has_s, self._has_t = s is not None, t is not None
        if s is None:
            self.s = lambda x: torch.zeros_like(x)
        else:
            self.add_module("s", s)
        if t is None:
            self.t = lambda x: torch.zeros_like(x)
        else:
            self.add_module("t", t)

    def _run(self, z, inverse):
        dt = torch.promote_types(self.b.dtype, z.dtype)      # `self.b * z` promotes (coupling.py:203)
        mask, z = self.b.to(dt), z.to(dt)
        z_masked = mask * z
        scale = self.s(z_masked) if self._has_s else None
        trans = self.t(z_masked) if self._has_t else None
        return ops.masked_affine(z, mask.reshape(-1).contiguous(), scale, trans, inverse)

    def forward(self, z):
        return self._run(z, False)

    def inverse(self, z):
        return self._run(z, True)


class AffineCouplingBlock(Flow):
    """Split -> AffineCoupling -> Merge (coupling.py:225-259) as ONE kernel after the conditioner:
    `vcnf_affine_coupling_fwd` copies the conditioning half, transforms the other half and reduces
    the per-sample log|det|.  The three sub-modules are kept so `state_dict` keys match
    (`flows.1.param_map...`)."""

    def __init__(self, param_map, scale=True, scale_map="exp", split_mode="channel"):
        super().__init__()
        self.flows = nn.ModuleList([Split(split_mode), AffineCoupling(param_map, scale, scale_map),
                                    Merge(split_mode)])
        self.split_mode = split_mode

    def _run(self, z, inverse):
        if "checkerboard" in self.split_mode:
            # checkerboard split (reshape.py:30-44): the three layers in sequence, index-map kernels either side
            log_det_tot = torch.zeros(z.shape[0], dtype=z.dtype, device=z.device)
            for flow in (reversed(self.flows) if inverse else self.flows):
                z, log_det = flow.inverse(z) if inverse else flow(z)
                log_det_tot = log_det_tot + log_det
            return z, log_det_tot
        if self.split_mode not in ("channel", "channel_inv"):
            raise NotImplementedError("Mode " + self.split_mode + " is not implemented.")
        coupling = self.flows[1]
        C = z.shape[1]
        first = (C + 1) // 2                                   # chunk(2): first half = ceil(C/2)
        if self.split_mode == "channel":
            z1, c2_start, C2 = z[:, :first], first, C - first
        else:
            z1, c2_start, C2 = z[:, first:], 0, first
        param = coupling.param_map(z1)
        out, ld = ops.affine_coupling(z, param, c2_start, C2, coupling._map_code(), inverse)
        if not coupling.scale:
            ld = torch.zeros(z.shape[0], dtype=z.dtype, device=z.device)
        return out, ld

    def forward(self, z):
        return self._run(z, False)

    def inverse(self, z):
        return self._run(z, True)
