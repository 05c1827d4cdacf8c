"""Glow block (reference: normflow/flows/affine/glow.py:12-75)."""
import torch
from torch import nn

from ..base import Flow
from .coupling import AffineCouplingBlock
from ..mixing import Invertible1x1Conv
from ..normalization import ActNorm
from ... import nets, ops


class GlowBlock(Flow):
    """AffineCouplingBlock(ConvNet2d 3x3,1x1,3x3) + Invertible1x1Conv + ActNorm."""

    def __init__(self, channels, hidden_channels, scale=True, scale_map="sigmoid",
                 split_mode="channel", leaky=0.0, init_zeros=True, use_lu=True,
                 net_actnorm=False):
        super().__init__()
        self.flows = nn.ModuleList([])
        kernel_size = (3, 1, 3)
        num_param = 2 if scale else 1
        if "channel" == split_mode:
            channels_ = (channels // 2,) + 2 * (hidden_channels,)
            channels_ += (num_param * ((channels + 1) // 2),)
        elif "channel_inv" == split_mode:
            channels_ = ((channels + 1) // 2,) + 2 * (hidden_channels,)
            channels_ += (num_param * (channels // 2),)
        else:
            raise NotImplementedError("Mode " + split_mode + " is not implemented.")
        param_map = nets.ConvNet2d(channels_, kernel_size, leaky, init_zeros, actnorm=net_actnorm)
        self.flows += [AffineCouplingBlock(param_map, scale, scale_map, split_mode)]
        if channels > 1:
            self.flows += [Invertible1x1Conv(channels, use_lu)]
        self.flows += [ActNorm((channels,) + (1, 1))]

    @staticmethod
    def _total(log_dets, z):
        """Sum of the sub-layers' log-dets: per-sample vectors and 0-dim constants are added separately and combined once
        (same value as the reference's running `log_det_tot += log_det` up to summation order; one broadcast add and one
        batch reduction in the backward instead of one per constant)."""
        vec = [ld for ld in log_dets if torch.is_tensor(ld) and ld.dim() > 0]
        const = [ld for ld in log_dets if not (torch.is_tensor(ld) and ld.dim() > 0)]
        tot = None
        for ld in vec:
            tot = ld if tot is None else tot + ld
        c = None
        for ld in const:
            c = ld if c is None else c + ld
        if tot is None:
            tot = torch.zeros(z.shape[0], dtype=z.dtype, device=z.device)
        return tot if c is None else tot + c

    def forward(self, z):
        lds = []
        for flow in self.flows:
            z, log_det = flow(z)
            lds.append(log_det)
        return z, self._total(lds, z)

    def _fused_tail(self, z):
        """ActNorm.inverse + Invertible1x1Conv.inverse as one node (ops.glow_actnorm_conv) when the block has the
        LU-parametrised convolution, the ActNorm is initialised and everything is one CUDA dtype; None otherwise."""
        if len(self.flows) != 3:
            return None
        conv, an = self.flows[1], self.flows[2]
        if not (isinstance(conv, Invertible1x1Conv) and conv.use_lu and isinstance(an, ActNorm)):
            return None
        if not (z.is_cuda and z.dim() == 4 and conv.num_channels <= ops.GLOW_W_MAX and an.initialised()):
            return None
        if not (z.dtype == conv.L.dtype == conv.P.dtype == an.s.dtype == an.t.dtype) or an.s.numel() != z.shape[1]:
            return None
        if not (isinstance(an.s, nn.Parameter) and isinstance(an.t, nn.Parameter)):
            return None
        return ops.glow_actnorm_conv(z, conv.L, conv.U, conv.log_S, an.s, an.t, conv.P, conv.sign_S)

    def inverse(self, z):
        lds = []
        fused = self._fused_tail(z)
        if fused is not None:
            z, ld = fused
            z, ld_c = self.flows[0].inverse(z)
            return z, self._total([ld, ld_c], z)
        for i in range(len(self.flows) - 1, -1, -1):
            z, log_det = self.flows[i].inverse(z)
            lds.append(log_det)
        return z, self._total(lds, z)
