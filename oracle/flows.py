"""Oracle bijector layers (plain torch, CPU).  API: `forward(z) -> (z', log_det)`,
`inverse(z) -> (z', log_det)` as `normflow/flows/base.py:6-21`.  Module trees mirror the
reference `state_dict` layout (SURVEY.md 8b "Checkpoint keys").
"""
import math

import numpy as np
import torch
from torch import nn
import torch.nn.functional as F

from . import nets, splines


class Flow(nn.Module):
    def forward(self, z):
        raise NotImplementedError("Forward pass has not been implemented.")

    def inverse(self, z):
        raise NotImplementedError("This flow has no algebraic inverse.")


class Reverse(Flow):
    """flows/base.py:24-40."""

    def __init__(self, flow):
        super().__init__()
        self.flow = flow

    def forward(self, z):
        return self.flow.inverse(z)

    def inverse(self, z):
        return self.flow.forward(z)


# ----------------------------------------------------------------------------- affine family
class AffineConstFlow(Flow):
    """flows/affine/coupling.py:10-53: z*exp(s)+t, log_det = prod(spatial)*sum(s) (0-dim)."""

    def __init__(self, shape, scale=True, shift=True):
        super().__init__()
        shape = (shape,) if isinstance(shape, int) else tuple(shape)
        zeros = torch.zeros(shape)[None]
        if scale:
            self.s = nn.Parameter(zeros.clone())
        else:
            self.register_buffer("s", zeros.clone())
        if shift:
            self.t = nn.Parameter(zeros.clone())
        else:
            self.register_buffer("t", zeros.clone())
        self.batch_dims = [i for i, n in enumerate(self.s.shape) if n == 1]

    def _spatial(self, z):
        return int(np.prod([z.size(i) for i in self.batch_dims[1:]])) if len(self.batch_dims) > 1 else 1

    def forward(self, z):
        return z * torch.exp(self.s) + self.t, self._spatial(z) * torch.sum(self.s)

    def inverse(self, z):
        return (z - self.t) * torch.exp(-self.s), -self._spatial(z) * torch.sum(self.s)


class ActNorm(AffineConstFlow):
    """flows/normalization.py:8-38: data-dependent init on the first call, either direction."""

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        self.register_buffer("data_dep_init_done", torch.tensor(0.0))

    def forward(self, z):
        if not self.data_dep_init_done > 0.0:
            s0 = -torch.log(z.std(dim=self.batch_dims, keepdim=True) + 1e-6)
            self.s.data = s0.data
            self.t.data = (-z.mean(dim=self.batch_dims, keepdim=True) * torch.exp(self.s)).data
            self.data_dep_init_done = torch.tensor(1.0)
        return super().forward(z)

    def inverse(self, z):
        if not self.data_dep_init_done:
            self.s.data = torch.log(z.std(dim=self.batch_dims, keepdim=True) + 1e-6).data
            self.t.data = z.mean(dim=self.batch_dims, keepdim=True).data
            self.data_dep_init_done = torch.tensor(1.0)
        return super().inverse(z)


class Split(Flow):
    """flows/reshape.py:9-55, modes channel / channel_inv (first chunk = ceil(C/2))."""

    def __init__(self, mode="channel"):
        super().__init__()
        if mode not in ("channel", "channel_inv"):
            raise NotImplementedError("Mode " + mode + " is not implemented.")
        self.mode = mode

    def forward(self, z):
        a, b = z.chunk(2, dim=1)
        return ([a, b] if self.mode == "channel" else [b, a]), 0

    def inverse(self, z):
        z1, z2 = z
        return torch.cat([z1, z2] if self.mode == "channel" else [z2, z1], 1), 0


class Merge(Split):
    def forward(self, z):
        return Split.inverse(self, z)

    def inverse(self, z):
        return Split.forward(self, z)


class Squeeze(Flow):
    """flows/reshape.py:93-117: forward out[b,c,2h+i,2w+j] = in[b,4c+2i+j,h,w]."""

    def forward(self, z):
        b, c, h, w = z.shape
        z = z.view(b, c // 4, 2, 2, h, w).permute(0, 1, 4, 2, 5, 3).contiguous()
        return z.view(b, c // 4, 2 * h, 2 * w), 0

    def inverse(self, z):
        b, c, h, w = z.shape
        z = z.view(b, c, h // 2, 2, w // 2, 2).permute(0, 1, 3, 5, 2, 4).contiguous()
        return z.view(b, 4 * c, h // 2, w // 2), 0


class AffineCoupling(Flow):
    """flows/affine/coupling.py:94-168.  shift = param[:,0::2], scale = param[:,1::2]."""

    def __init__(self, param_map, scale=True, scale_map="exp"):
        super().__init__()
        self.add_module("param_map", param_map)
        self.scale, self.scale_map = scale, scale_map

    def forward(self, z):
        z1, z2 = z
        p = self.param_map(z1)
        if not self.scale:
            return [z1, z2 + p], 0
        shift, raw = p[:, 0::2, ...], p[:, 1::2, ...]
        dims = list(range(1, shift.dim()))
        if self.scale_map == "exp":
            return [z1, z2 * torch.exp(raw) + shift], torch.sum(raw, dim=dims)
        sc = torch.sigmoid(raw + 2)
        if self.scale_map == "sigmoid":
            return [z1, z2 / sc + shift], -torch.sum(torch.log(sc), dim=dims)
        if self.scale_map == "sigmoid_inv":
            return [z1, z2 * sc + shift], torch.sum(torch.log(sc), dim=dims)
        raise NotImplementedError("This scale map is not implemented.")

    def inverse(self, z):
        z1, z2 = z
        p = self.param_map(z1)
        if not self.scale:
            return [z1, z2 - p], 0
        shift, raw = p[:, 0::2, ...], p[:, 1::2, ...]
        dims = list(range(1, shift.dim()))
        if self.scale_map == "exp":
            return [z1, (z2 - shift) * torch.exp(-raw)], -torch.sum(raw, dim=dims)
        sc = torch.sigmoid(raw + 2)
        if self.scale_map == "sigmoid":
            return [z1, (z2 - shift) * sc], torch.sum(torch.log(sc), dim=dims)
        if self.scale_map == "sigmoid_inv":
            return [z1, (z2 - shift) / sc], -torch.sum(torch.log(sc), dim=dims)
        raise NotImplementedError("This scale map is not implemented.")


class _Chain(Flow):
    """Shared loop of AffineCouplingBlock (coupling.py:247-259) and GlowBlock (glow.py:63-75)."""

    def forward(self, z):
        tot = torch.zeros(z.shape[0], dtype=z.dtype, device=z.device)
        for f in self.flows:
            z, ld = f(z)
            tot = tot + ld
        return z, tot

    def inverse(self, z):
        tot = torch.zeros(z.shape[0], dtype=z.dtype, device=z.device)
        for f in reversed(self.flows):
            z, ld = f.inverse(z)
            tot = tot + ld
        return z, tot


class AffineCouplingBlock(_Chain):
    def __init__(self, param_map, scale=True, scale_map="exp", split_mode="channel"):
        super().__init__()
        self.flows = nn.ModuleList([Split(split_mode), AffineCoupling(param_map, scale, scale_map),
                                    Merge(split_mode)])


class MaskedAffineFlow(Flow):
    """flows/affine/coupling.py:171-222: b*z + (1-b)*(z*exp(s(b*z)) + t(b*z)); +-inf -> NaN."""

    def __init__(self, b, t=None, s=None):
        super().__init__()
        self.register_buffer("b", b.view(1, *b.size()))
        self.s = s if s is not None else (lambda x: torch.zeros_like(x))
        self.t = t if t is not None else (lambda x: torch.zeros_like(x))

    def _st(self, zm, z):
        nan = torch.tensor(np.nan, dtype=z.dtype, device=z.device)
        sc = self.s(zm)
        sc = torch.where(torch.isfinite(sc), sc, nan)
        tr = self.t(zm)
        tr = torch.where(torch.isfinite(tr), tr, nan)
        return sc, tr

    def forward(self, z):
        zm = self.b * z
        sc, tr = self._st(zm, z)
        out = zm + (1 - self.b) * (z * torch.exp(sc) + tr)
        return out, torch.sum((1 - self.b) * sc, dim=list(range(1, self.b.dim())))

    def inverse(self, z):
        zm = self.b * z
        sc, tr = self._st(zm, z)
        out = zm + (1 - self.b) * (z - tr) * torch.exp(-sc)
        return out, -torch.sum((1 - self.b) * sc, dim=list(range(1, self.b.dim())))


# ----------------------------------------------------------------------------- mixing
class Permute(Flow):
    """flows/mixing.py:10-54.  swap: forward splits at C//2, inverse at (C+1)//2."""

    def __init__(self, num_channels, mode="shuffle"):
        super().__init__()
        self.mode, self.num_channels = mode, num_channels
        if mode == "shuffle":
            perm = torch.randperm(num_channels)
            self.register_buffer("perm", perm)
            self.register_buffer("inv_perm", torch.argsort(perm))

    def _swap(self, z, cut):
        return torch.cat([z[:, cut:, ...], z[:, :cut, ...]], dim=1)

    def forward(self, z):
        if self.mode == "shuffle":
            return z[:, self.perm, ...], 0
        if self.mode == "swap":
            return self._swap(z, self.num_channels // 2), 0
        raise NotImplementedError("The mode " + self.mode + " is not implemented.")

    def inverse(self, z):
        if self.mode == "shuffle":
            return z[:, self.inv_perm, ...], 0
        if self.mode == "swap":
            return self._swap(z, (self.num_channels + 1) // 2), 0
        raise NotImplementedError("The mode " + self.mode + " is not implemented.")


class Invertible1x1Conv(Flow):
    """flows/mixing.py:57-128.  inverse (density dir): z' = W z per pixel, W = P L U;
    forward uses W^-1 = U^-1 L^-1 P^T with the triangular inverses taken in fp64."""

    def __init__(self, num_channels, use_lu=False):
        super().__init__()
        self.num_channels, self.use_lu = num_channels, use_lu
        q = torch.linalg.qr(torch.randn(num_channels, num_channels))[0]
        if use_lu:
            p, l, u = torch.linalg.lu(q)
            self.register_buffer("P", p)
            self.L = nn.Parameter(l)
            s = u.diag()
            self.register_buffer("sign_S", torch.sign(s))
            self.log_S = nn.Parameter(torch.log(torch.abs(s)))
            self.U = nn.Parameter(torch.triu(u, diagonal=1))
            self.register_buffer("eye", torch.diag(torch.ones(num_channels)))
        else:
            self.W = nn.Parameter(q)

    def _matrix(self, inverse):
        if not self.use_lu:
            if not inverse:
                return self.W, torch.slogdet(self.W)[1]
            wi = torch.inverse(self.W.double()).type(self.W.dtype)
            return wi, -torch.slogdet(self.W)[1]
        low = torch.tril(self.L, diagonal=-1) + self.eye
        up = torch.triu(self.U, diagonal=1) + torch.diag(self.sign_S * torch.exp(self.log_S))
        if not inverse:
            return self.P @ low @ up, torch.sum(self.log_S)
        dt = self.log_S.dtype
        li = torch.inverse(low.double()).type(dt)
        ui = torch.inverse(up.double()).type(dt)
        return ui @ li @ self.P.t(), -torch.sum(self.log_S)

    def _conv(self, z, inverse):
        w, ld = self._matrix(inverse)
        c = self.num_channels
        return F.conv2d(z, w.view(c, c, 1, 1)), ld * z.size(2) * z.size(3)

    def forward(self, z):
        return self._conv(z, True)

    def inverse(self, z):
        return self._conv(z, False)


class _Permutation(nn.Module):
    def __init__(self, permutation):
        super().__init__()
        self.register_buffer("_permutation", permutation)


class _LULinear(nn.Module):
    """flows/mixing.py:352-462: W = L U, L unit-lower, diag(U) = softplus(d) + eps."""

    def __init__(self, features, identity_init=True, eps=1e-3):
        super().__init__()
        self.features, self.eps = features, eps
        n_tri = ((features - 1) * features) // 2
        self.bias = nn.Parameter(torch.zeros(features))
        self.lower_entries = nn.Parameter(torch.zeros(n_tri))
        self.upper_entries = nn.Parameter(torch.zeros(n_tri))
        self.unconstrained_upper_diag = nn.Parameter(torch.zeros(features))
        if identity_init:
            nn.init.constant_(self.unconstrained_upper_diag, math.log(math.exp(1 - eps) - 1))
        else:
            sd = 1.0 / math.sqrt(features)
            for p in (self.lower_entries, self.upper_entries, self.unconstrained_upper_diag):
                nn.init.uniform_(p, -sd, sd)

    def lower_upper(self):
        d = self.features
        li = np.tril_indices(d, k=-1)
        ui = np.triu_indices(d, k=1)
        di = np.diag_indices(d)
        low = self.lower_entries.new_zeros(d, d)
        low[li[0], li[1]] = self.lower_entries
        low[di[0], di[1]] = 1.0
        up = self.upper_entries.new_zeros(d, d)
        up[ui[0], ui[1]] = self.upper_entries
        up[di[0], di[1]] = F.softplus(self.unconstrained_upper_diag) + self.eps
        return low, up

    def logabsdet(self):
        return torch.sum(torch.log(F.softplus(self.unconstrained_upper_diag) + self.eps))


class LULinearPermute(Flow):
    """flows/mixing.py:465-492.  inverse (density dir): z[:,perm] -> .U^T -> .L^T + bias;
    forward: subtract bias, unit-lower solve, upper solve, inverse permutation."""

    def __init__(self, num_channels, identity_init=True):
        super().__init__()
        self.permutation = _Permutation(torch.randperm(num_channels))
        self.linear = _LULinear(num_channels, identity_init=identity_init)

    def forward(self, z):
        low, up = self.linear.lower_upper()
        y = (z - self.linear.bias).t()
        y = torch.linalg.solve_triangular(low, y, upper=False, unitriangular=True)
        y = torch.linalg.solve_triangular(up, y, upper=True).t()
        y = y[:, torch.argsort(self.permutation._permutation)]
        return y, (-self.linear.logabsdet() * z.new_ones(z.shape[0])).view(-1)

    def inverse(self, z):
        low, up = self.linear.lower_upper()
        y = z[:, self.permutation._permutation]
        y = F.linear(F.linear(y, up), low, self.linear.bias)
        return y, (self.linear.logabsdet() * z.new_ones(z.shape[0])).view(-1)


class GlowBlock(_Chain):
    """flows/affine/glow.py:12-75: coupling(ConvNet2d 3x3,1x1,3x3) + 1x1 conv + ActNorm."""

    def __init__(self, channels, hidden_channels, scale=True, scale_map="sigmoid",
                 split_mode="channel", leaky=0.0, init_zeros=True, use_lu=True):
        super().__init__()
        n_par = 2 if scale else 1
        if split_mode == "channel":
            ch = (channels // 2, hidden_channels, hidden_channels, n_par * ((channels + 1) // 2))
        elif split_mode == "channel_inv":
            ch = ((channels + 1) // 2, hidden_channels, hidden_channels, n_par * (channels // 2))
        else:
            raise NotImplementedError("Mode " + split_mode + " is not implemented.")
        net = nets.ConvNet2d(ch, (3, 1, 3), leaky, init_zeros)
        mods = [AffineCouplingBlock(net, scale, scale_map, split_mode)]
        if channels > 1:
            mods.append(Invertible1x1Conv(channels, use_lu))
        mods.append(ActNorm((channels, 1, 1)))
        self.flows = nn.ModuleList(mods)


# ----------------------------------------------------------------------------- autoregressive
class _Autoregressive(Flow):
    """flows/affine/autoregressive.py:13-36: forward = one net pass; inverse = D passes."""

    def forward(self, x, context=None):
        return self._elementwise(x, self.autoregressive_net(x), inverse=False)

    def inverse(self, x, context=None):
        out = torch.zeros_like(x)
        lad = None
        for _ in range(int(np.prod(x.shape[1:]))):
            out, lad = self._elementwise(x, self.autoregressive_net(out), inverse=True)
        return out, lad


class MaskedAffineAutoregressive(_Autoregressive):
    """flows/affine/autoregressive.py:48-103: scale = sigmoid(u+2)+1e-3; params [B,D,2]."""

    def __init__(self, features, hidden_features, num_blocks=2, activation=F.relu):
        super().__init__()
        self.features = features
        self.autoregressive_net = nets.MADE(features, hidden_features, num_blocks=num_blocks,
                                            output_multiplier=2, activation=activation)

    def _elementwise(self, x, params, inverse):
        p = params.view(-1, self.features, 2)
        scale = torch.sigmoid(p[..., 0] + 2.0) + 1e-3
        log_scale = torch.log(scale)
        if inverse:
            return (x - p[..., 1]) / scale, -log_scale.sum(-1)
        return scale * x + p[..., 1], log_scale.sum(-1)


# ----------------------------------------------------------------------------- neural spline flows
def _params_per_feature(tails, k):
    """neural_spline/coupling.py:301-307 == autoregressive.py:84-90."""
    return {"linear": 3 * k - 1, "circular": 3 * k}.get(tails if isinstance(tails, str) else None, 3 * k + 1)


def alternating_mask(features, even):
    """utils/masks.py:5-15."""
    m = torch.zeros(features, dtype=torch.uint8)
    m[(0 if even else 1)::2] += 1
    return m


class _SharedSpline(Flow):
    """`PiecewiseRationalQuadraticCDF`, neural_spline/coupling.py:165-246: params shared over
    the batch ([F,K],[F,K],[F,K-1|K|K+1])."""

    def __init__(self, features, num_bins, tails, tail_bound):
        super().__init__()
        self.tails = tails
        if torch.is_tensor(tail_bound):
            self.register_buffer("tail_bound", tail_bound)
        else:
            self.tail_bound = tail_bound
        nd = _params_per_feature(tails, num_bins) - 2 * num_bins
        self.unnormalized_widths = nn.Parameter(torch.zeros(features, num_bins))
        self.unnormalized_heights = nn.Parameter(torch.zeros(features, num_bins))
        self.unnormalized_derivatives = nn.Parameter(
            splines.boundary_derivative_constant() * torch.ones(features, nd))

    def _run(self, x, inverse):
        n = x.shape[0]
        ex = lambda p: p[None].expand(n, *p.shape)
        y, lad = splines.unconstrained_rqs(
            x, ex(self.unnormalized_widths), ex(self.unnormalized_heights),
            ex(self.unnormalized_derivatives), inverse=inverse, tails=self.tails,
            tail_bound=self.tail_bound)
        return y, lad.sum(-1)

    def forward(self, x):
        return self._run(x, False)

    def inverse(self, x):
        return self._run(x, True)


class _SplineCoupling(Flow):
    """`PiecewiseRationalQuadraticCoupling` + `Coupling`, neural_spline/coupling.py:18-137,
    249-343, 2-D inputs, apply_unconditional_transform=True (wrapper.py:80)."""

    def __init__(self, mask, net_fn, num_bins, tails, tail_bound):
        super().__init__()
        mask = torch.as_tensor(mask)
        idx = torch.arange(len(mask))
        self.register_buffer("identity_features", idx.masked_select(mask <= 0))
        self.register_buffer("transform_features", idx.masked_select(mask > 0))
        self.num_bins = num_bins
        if isinstance(tails, (list, tuple)):
            self.tails = [tails[i] for i in self.transform_features]
            id_tails = [tails[i] for i in self.identity_features]
        else:
            self.tails = id_tails = tails
        if torch.is_tensor(tail_bound):
            id_bound = tail_bound[self.identity_features]
        else:
            self.tail_bound = id_bound = tail_bound
        self.p = _params_per_feature(self.tails, num_bins)
        self.transform_net = net_fn(len(self.identity_features), len(self.transform_features) * self.p)
        self.unconditional_transform = _SharedSpline(len(self.identity_features), num_bins,
                                                     id_tails, id_bound)
        if torch.is_tensor(tail_bound):
            self.register_buffer("tail_bound", tail_bound[self.transform_features])

    def _conditional(self, x, params, inverse):
        k = self.num_bins
        p = params.reshape(x.shape[0], x.shape[1], -1)
        root_h = np.sqrt(self.transform_net.hidden_features)         # coupling.py:314-316
        uw, uh, ud = p[..., :k] / root_h, p[..., k:2 * k] / root_h, p[..., 2 * k:]
        y, lad = splines.unconstrained_rqs(x, uw, uh, ud, inverse=inverse, tails=self.tails,
                                           tail_bound=self.tail_bound)
        return y, lad.sum(-1)

    def forward(self, x):                                            # coupling.py:70-96
        ident, trans = x[:, self.identity_features], x[:, self.transform_features]
        trans, lad = self._conditional(trans, self.transform_net(ident), False)
        ident, lad_id = self.unconditional_transform(ident)
        out = torch.empty_like(x)
        out[:, self.identity_features] = ident
        out[:, self.transform_features] = trans
        return out, lad + lad_id

    def inverse(self, x):                                            # coupling.py:98-125
        ident, trans = x[:, self.identity_features], x[:, self.transform_features]
        ident, lad = self.unconditional_transform.inverse(ident)
        trans, lad_t = self._conditional(trans, self.transform_net(ident), True)
        out = torch.empty_like(x)
        out[:, self.identity_features] = ident
        out[:, self.transform_features] = trans
        return out, lad + lad_t


class CoupledRationalQuadraticSpline(Flow):
    """neural_spline/wrapper.py:15-89 -- note the direction swap (:83-89)."""

    def __init__(self, num_input_channels, num_blocks, num_hidden_channels, num_bins=8,
                 tails="linear", tail_bound=3.0, activation=nn.ReLU, reverse_mask=False):
        super().__init__()
        net_fn = lambda i, o: nets.ResidualNet(i, o, num_hidden_channels, num_blocks, activation())
        self.prqct = _SplineCoupling(alternating_mask(num_input_channels, even=reverse_mask),
                                     net_fn, num_bins, tails, tail_bound)

    def forward(self, z):
        z, ld = self.prqct.inverse(z)
        return z, ld.view(-1)

    def inverse(self, z):
        z, ld = self.prqct(z)
        return z, ld.view(-1)


class CircularCoupledRationalQuadraticSpline(CoupledRationalQuadraticSpline):
    """neural_spline/wrapper.py:92-194: per-feature tails list + PeriodicFeatures on the
    circular identity coordinates; final layer initialised to the identity spline."""

    def __init__(self, num_input_channels, num_blocks, num_hidden_channels, ind_circ,
                 num_bins=8, tail_bound=3.0, activation=nn.ReLU, reverse_mask=False,
                 mask=None, init_identity=True):
        Flow.__init__(self)
        if mask is None:
            mask = alternating_mask(num_input_channels, even=reverse_mask)
        ident = torch.arange(num_input_channels).masked_select(torch.as_tensor(mask) <= 0)
        circ_id = [i for i, f in enumerate(ident.tolist()) if f in list(ind_circ)]
        scale = math.pi / tail_bound[circ_id] if torch.is_tensor(tail_bound) else math.pi / tail_bound

        def net_fn(i, o):
            pf = nets.PeriodicFeatures(i, circ_id, scale) if circ_id else None
            net = nets.ResidualNet(i, o, num_hidden_channels, num_blocks, activation(), preprocessing=pf)
            if init_identity:
                nn.init.constant_(net.final_layer.weight, 0.0)
                nn.init.constant_(net.final_layer.bias, splines.boundary_derivative_constant())
            return net

        tails = ["circular" if i in list(ind_circ) else "linear" for i in range(num_input_channels)]
        self.prqct = _SplineCoupling(mask, net_fn, num_bins, tails, tail_bound)


class _SplineAutoregressive(_Autoregressive):
    """`MaskedPiecewiseRationalQuadraticAutoregressive`, neural_spline/autoregressive.py:18-137.
    The MADE output is NOT divided by sqrt(hidden) (MADE lacks `hidden_features`, hazard A.8)."""

    def __init__(self, features, hidden_features, num_bins, tails, tail_bound, num_blocks,
                 permute_mask, activation, init_identity):
        super().__init__()
        self.num_bins, self.tails = num_bins, tails
        pre = None
        if isinstance(tails, (list, tuple)):
            circ = [i for i in range(features) if tails[i] == "circular"]
            scale = math.pi / tail_bound[circ] if torch.is_tensor(tail_bound) else math.pi / tail_bound
            pre = nets.PeriodicFeatures(features, circ, scale)
        self.p = _params_per_feature(tails, num_bins)
        self.autoregressive_net = nets.MADE(features, hidden_features, num_blocks=num_blocks,
                                            output_multiplier=self.p, permute_mask=permute_mask,
                                            activation=activation, preprocessing=pre)
        if init_identity:
            nn.init.constant_(self.autoregressive_net.final_layer.weight, 0.0)
            nn.init.constant_(self.autoregressive_net.final_layer.bias,
                              splines.boundary_derivative_constant())
        if torch.is_tensor(tail_bound):
            self.register_buffer("tail_bound", tail_bound)
        else:
            self.tail_bound = tail_bound

    def _elementwise(self, x, params, inverse):
        k = self.num_bins
        p = params.view(x.shape[0], x.shape[1], self.p)
        y, lad = splines.unconstrained_rqs(x, p[..., :k], p[..., k:2 * k], p[..., 2 * k:],
                                           inverse=inverse, tails=self.tails, tail_bound=self.tail_bound)
        return y, lad.sum(-1)


class AutoregressiveRationalQuadraticSpline(Flow):
    """neural_spline/wrapper.py:197-259 (direction swap at :253-259)."""

    def __init__(self, num_input_channels, num_blocks, num_hidden_channels, num_bins=8,
                 tail_bound=3, activation=nn.ReLU, permute_mask=False, init_identity=True):
        super().__init__()
        self.mprqat = _SplineAutoregressive(num_input_channels, num_hidden_channels, num_bins,
                                            "linear", tail_bound, num_blocks, permute_mask,
                                            activation(), init_identity)

    def forward(self, z):
        z, ld = self.mprqat.inverse(z)
        return z, ld.view(-1)

    def inverse(self, z):
        z, ld = self.mprqat(z)
        return z, ld.view(-1)


class CircularAutoregressiveRationalQuadraticSpline(AutoregressiveRationalQuadraticSpline):
    """neural_spline/wrapper.py:262-330."""

    def __init__(self, num_input_channels, num_blocks, num_hidden_channels, ind_circ, num_bins=8,
                 tail_bound=3, activation=nn.ReLU, permute_mask=True, init_identity=True):
        Flow.__init__(self)
        tails = ["circular" if i in list(ind_circ) else "linear" for i in range(num_input_channels)]
        self.mprqat = _SplineAutoregressive(num_input_channels, num_hidden_channels, num_bins,
                                            tails, tail_bound, num_blocks, permute_mask,
                                            activation(), init_identity)
