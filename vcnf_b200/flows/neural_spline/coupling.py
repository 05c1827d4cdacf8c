"""Neural-spline coupling layers (reference: normflow/flows/neural_spline/coupling.py).

`Coupling.forward` / `.inverse` run as: one column gather for the conditioner input, the
conditioner, then ONE `vcnf_rqs_layer_fwd` launch that does the conditional spline on the
transform features, the unconditional spline on the identity features, the scatter back to the
original columns and the per-sample log|det| sum (coupling.py:70-125,147-159,211-240,309-343)."""
import warnings

import numpy as np
import torch
from torch import nn

from ..base import Flow
from ... import ops
from ...utils import splines

_TAIL_CODE = {"linear": ops.TAIL_LINEAR, "circular": ops.TAIL_CIRCULAR}


def _params_per_feature(tails, num_bins):
    if tails == "linear":
        return num_bins * 3 - 1
    if tails == "circular":
        return num_bins * 3
    return num_bins * 3 + 1


def check_unit_box(values):
    """tails=None: the spline is defined on [0, 1] only and the reference refuses anything else
    (utils/splines.py:105-106, a host-side min / max like there)."""
    if values.numel() and (float(values.min()) < 0.0 or float(values.max()) > 1.0):
        raise ValueError("Input outside domain")


def _modes_and_bounds(tails, tail_bound, n):
    """Per-feature (tail mode, bound) lists for a segment of n features."""
    if tails is None:
        # bounded spline (neural_spline/coupling.py:321-323): `rational_quadratic_spline` is called with its default
        # box [0, 1] and `tail_bound` is not passed on; the kernels' TAIL_NONE interval is [0, 2 bound]
        return [ops.TAIL_NONE] * n, [0.5] * n
    if isinstance(tails, (list, tuple)):
        modes = [ops.TAIL_LIST_LINEAR if t == "linear" else ops.TAIL_LIST_CIRCULAR for t in tails]
    elif tails in _TAIL_CODE:
        modes = [_TAIL_CODE[tails]] * n
    else:
        raise RuntimeError("{} tails are not implemented.".format(tails))
    if torch.is_tensor(tail_bound):
        bounds = [float(b) for b in tail_bound.detach().cpu().reshape(-1)]
        if len(bounds) == 1:
            bounds = bounds * n
    else:
        bounds = [float(tail_bound)] * n
    return modes, bounds


class PiecewiseRationalQuadraticCDF(Flow):
    """Elementwise spline with parameters shared across the batch (coupling.py:165-246)."""

    def __init__(self, shape, num_bins=10, tails=None, tail_bound=1., identity_init=True,
                 min_bin_width=splines.DEFAULT_MIN_BIN_WIDTH,
                 min_bin_height=splines.DEFAULT_MIN_BIN_HEIGHT,
                 min_derivative=splines.DEFAULT_MIN_DERIVATIVE):
        super().__init__()
        splines.check_minimums(num_bins, min_bin_width, min_bin_height, min_derivative)
        if len(shape) not in (1, 3):
            raise NotImplementedError("spline parameters must be per feature [D] or per pixel [C, H, W]")
        self._img = len(shape) == 3
        if torch.is_tensor(tail_bound):
            self.register_buffer("tail_bound", tail_bound)
        else:
            self.tail_bound = tail_bound
        self.tails = tails
        self.num_bins = num_bins
        nd = _params_per_feature(tails, num_bins) - 2 * num_bins
        if identity_init:
            self.unnormalized_widths = nn.Parameter(torch.zeros(*shape, num_bins))
            self.unnormalized_heights = nn.Parameter(torch.zeros(*shape, num_bins))
            constant = np.log(np.exp(1 - min_derivative) - 1)
            self.unnormalized_derivatives = nn.Parameter(constant * torch.ones(*shape, nd))
        else:
            self.unnormalized_widths = nn.Parameter(torch.rand(*shape, num_bins))
            self.unnormalized_heights = nn.Parameter(torch.rand(*shape, num_bins))
            self.unnormalized_derivatives = nn.Parameter(torch.rand(*shape, nd))
        self._spec = None


    def _load_from_state_dict(self, *args, **kwargs):
        # the kernel spec caches the tail bounds: a loaded `tail_bound` buffer must rebuild it
        super()._load_from_state_dict(*args, **kwargs)
        self._spec = None

    def packed_params(self):
        return torch.cat([self.unnormalized_widths, self.unnormalized_heights,
                          self.unnormalized_derivatives], dim=-1)

    def segment(self):
        n = self.unnormalized_widths.shape[0]
        return _modes_and_bounds(self.tails, self.tail_bound, n)

    def pixel_rows(self, batch):
        """Image-shaped parameters [C, H, W, P] laid out as the per-row parameters of NHWC pixel rows:
        [batch*H*W, C*P] (the batch expansion of coupling.py:207-216; its adjoint sums over the batch)."""
        p = self.packed_params()                                   # [C, H, W, P]
        C, H, W, P = p.shape
        rows = p.permute(1, 2, 0, 3).reshape(H * W, C * P)
        return rows.unsqueeze(0).expand(batch, H * W, C * P).reshape(batch * H * W, C * P)

    def _spline_image(self, inputs, inverse):
        B, C, H, W = inputs.shape
        if self._spec is None:
            modes, bounds = self.segment()
            P = _params_per_feature(self.tails, self.num_bins)
            cmap = [(ops.COL_COND << 16) | i for i in range(C)]
            self._spec = ops.SplineSpec(C, self.num_bins, cmap, modes, bounds, P, 1.0, [], [], P,
                                        isinstance(self.tails, (list, tuple)))
        rows = inputs.permute(0, 2, 3, 1).reshape(B * H * W, C)
        y, ld = ops.rqs_layer(rows, self.pixel_rows(B), None, self._spec, inverse)
        return y.reshape(B, H, W, C).permute(0, 3, 1, 2).contiguous(), ld.reshape(B, H * W).sum(1)

    def _spline(self, inputs, inverse=False):
        if self.tails is None:
            check_unit_box(inputs)
        if self._img:
            return self._spline_image(inputs, inverse)
        if self._spec is None:
            n = self.unnormalized_widths.shape[0]
            modes, bounds = self.segment()
            P = _params_per_feature(self.tails, self.num_bins)
            cmap = [(ops.COL_SHARED << 16) | i for i in range(n)]
            self._spec = ops.SplineSpec(n, self.num_bins, cmap, [], [], P, 1.0, modes, bounds, P,
                                        isinstance(self.tails, (list, tuple)))
        return ops.rqs_layer(inputs, None, self.packed_params(), self._spec, inverse)

    def forward(self, inputs, context=None):
        return self._spline(inputs, inverse=False)

    def inverse(self, inputs, context=None):
        return self._spline(inputs, inverse=True)


class Coupling(Flow):
    """Base coupling layer on a binary feature mask (coupling.py:18-137): 2-D inputs [B, D], or images [B, C, H, W]
    split along the channels (coupling.py:147-151), which run as NHWC pixel rows through the same spline kernel."""

    def __init__(self, mask, transform_net_create_fn, unconditional_transform=None):
        mask = torch.as_tensor(mask)
        if mask.dim() != 1:
            raise ValueError("Mask must be a 1-dim tensor.")
        if mask.numel() <= 0:
            raise ValueError("Mask can't be empty.")
        super().__init__()
        self.features = len(mask)
        features_vector = torch.arange(self.features)
        self.register_buffer("identity_features", features_vector.masked_select(mask <= 0))
        self.register_buffer("transform_features", features_vector.masked_select(mask > 0))
        assert self.num_identity_features + self.num_transform_features == self.features
        self.transform_net = transform_net_create_fn(
            self.num_identity_features,
            self.num_transform_features * self._transform_dim_multiplier())
        if unconditional_transform is None:
            self.unconditional_transform = None
        else:
            self.unconditional_transform = unconditional_transform(features=self.num_identity_features)
        # The library's own conditioner takes / returns rows with a padded (16-byte multiple) pitch, which
        # keeps every tile of the layer TMA-movable; any other user-supplied net sees plain contiguous tensors.
        self._pad_rows = bool(getattr(self.transform_net, "accepts_padded_rows", False))
        if self._pad_rows and hasattr(self.transform_net, "padded_out"):
            self.transform_net.padded_out = True

    @property
    def num_identity_features(self):
        return len(self.identity_features)

    @property
    def num_transform_features(self):
        return len(self.transform_features)

    def _check(self, inputs):
        if inputs.dim() not in [2, 4]:
            raise ValueError("Inputs must be a 2D or a 4D tensor.")
        if inputs.shape[1] != self.features:
            raise ValueError("Expected features = {}, got {}.".format(self.features, inputs.shape[1]))

    def _transform_dim_multiplier(self):
        raise NotImplementedError()


class PiecewiseRationalQuadraticCoupling(Coupling):
    """Rational-quadratic spline coupling (coupling.py:249-343)."""

    def __init__(self, mask, transform_net_create_fn, num_bins=10, tails=None, tail_bound=1.,
                 apply_unconditional_transform=False, img_shape=None,
                 min_bin_width=splines.DEFAULT_MIN_BIN_WIDTH,
                 min_bin_height=splines.DEFAULT_MIN_BIN_HEIGHT,
                 min_derivative=splines.DEFAULT_MIN_DERIVATIVE):
        splines.check_minimums(num_bins, min_bin_width, min_bin_height, min_derivative)
        self.num_bins = num_bins
        mask = torch.as_tensor(mask)
        features_vector = torch.arange(len(mask))
        identity_features = features_vector.masked_select(mask <= 0)
        transform_features = features_vector.masked_select(mask > 0)
        if isinstance(tails, (list, tuple)):
            self.tails = [tails[i] for i in transform_features]
            tails_ = [tails[i] for i in identity_features]
        else:
            self.tails = tails
            tails_ = tails
        if torch.is_tensor(tail_bound):
            tail_bound_ = tail_bound[identity_features]
        else:
            self.tail_bound = tail_bound
            tail_bound_ = tail_bound
        if apply_unconditional_transform:
            unconditional_transform = lambda features: PiecewiseRationalQuadraticCDF(
                shape=[features] + (img_shape if img_shape else []), num_bins=num_bins,
                tails=tails_, tail_bound=tail_bound_, min_bin_width=min_bin_width,
                min_bin_height=min_bin_height, min_derivative=min_derivative)
        else:
            unconditional_transform = None
        super().__init__(mask, transform_net_create_fn, unconditional_transform=unconditional_transform)
        if torch.is_tensor(tail_bound):
            self.register_buffer("tail_bound", tail_bound[transform_features])
        self._specs = None
        self._fused = None


    def _load_from_state_dict(self, *args, **kwargs):
        # the kernel specs cache the tail bounds: a loaded `tail_bound` buffer must rebuild them
        super()._load_from_state_dict(*args, **kwargs)
        self._specs = None
        self._fused = None

    def _transform_dim_multiplier(self):
        return _params_per_feature(self.tails, self.num_bins)

    def _shared_params(self):
        ut = self.unconditional_transform
        return ut.packed_params() if ut is not None else None

    def _needs_grad(self, inputs):
        return torch.is_grad_enabled() and (inputs.requires_grad or any(p.requires_grad for p in self.parameters()))

    def _fused_route(self, inputs, context):
        """A dense ResidualNet conditioner and the spline run as ONE kernel per layer (vcnf_resnet_spline_fwd).  Without
        gradients nothing but x and y touches HBM; with gradients (density direction) the same launch also writes what
        the backward kernels need (layer outputs, spline parameters) and the layer-by-layer autograd nodes are built
        around those values.  Returns the ops.FusedNetSpec, or None for a configuration the kernel does not cover."""
        net = self.transform_net
        ut = self.unconditional_transform
        if not ops.FUSED_NET:
            return None
        if (context is not None or (ut is not None and getattr(ut, "_img", False)) or inputs.dim() != 2
                or not inputs.is_cuda or inputs.dtype != torch.float32 or not hasattr(net, "fusable") or not net.fusable()):
            return None
        if self._fused is None:
            if self._specs is None:
                self._build_specs()
            trans = self.transform_features.tolist()
            cmodes, cbounds = _modes_and_bounds(self.tails, self.tail_bound, len(trans))
            smodes, sbounds = ut.segment() if ut is not None else ((), ())
            Ps = _params_per_feature(ut.tails, self.num_bins) if ut is not None else 0
            spec = ops.FusedNetSpec(self.identity_features.tolist(), trans, cmodes, cbounds, self.num_bins,
                                    self._transform_dim_multiplier(), self._wh_scale(),
                                    isinstance(self.tails, (list, tuple)), net.hidden_features, len(net.blocks),
                                    smodes, sbounds, Ps) \
                if self.features <= 32 and len(trans) <= 32 and len(self.identity_features) <= 32 else None
            self._fused = spec if spec is not None and spec.supported else False
        return self._fused or None

    def _check_domain(self, inputs):
        """tails=None: the values that reach a bounded spline must lie in [0, 1] (utils/splines.py:105-106)."""
        ut = self.unconditional_transform
        both = self.tails is None and ut is not None and ut.tails is None
        if both:
            check_unit_box(inputs)
        elif self.tails is None:
            check_unit_box(inputs.index_select(1, self.transform_features.to(inputs.device)))
        elif ut is not None and ut.tails is None:
            check_unit_box(inputs.index_select(1, self.identity_features.to(inputs.device)))

    def _wh_scale(self):
        net = self.transform_net
        if hasattr(net, "hidden_features"):
            return 1.0 / np.sqrt(net.hidden_features)           # coupling.py:314-316
        if hasattr(net, "hidden_channels"):
            return 1.0 / np.sqrt(net.hidden_channels)
        warnings.warn("Inputs to the softmax are not scaled down: initialization might be bad.")
        return 1.0

    def _build_specs(self):
        D, K = self.features, self.num_bins
        ident = self.identity_features.tolist()
        trans = self.transform_features.tolist()
        Pc = self._transform_dim_multiplier()
        cmodes, cbounds = _modes_and_bounds(self.tails, self.tail_bound, len(trans))
        list_mode = isinstance(self.tails, (list, tuple))
        ut = self.unconditional_transform
        img_ut = ut is not None and getattr(ut, "_img", False)
        if ut is not None:
            smodes, sbounds = ut.segment()
            Ps = _params_per_feature(ut.tails, K)
        else:
            smodes, sbounds, Ps = [], [], Pc
        if img_ut:
            # per-pixel parameters are not shared across rows: they enter as per-row parameters of their own launch
            n_id = len(ident)
            self._ident_rows_spec = ops.SplineSpec(n_id, K, [(ops.COL_COND << 16) | j for j in range(n_id)],
                                                   smodes, sbounds, Ps, 1.0, [], [], Ps,
                                                   isinstance(ut.tails, (list, tuple)))
            ut, smodes, sbounds = None, [], []
        wh = self._wh_scale()
        cond_map, shared_map, both = [0] * D, [0] * D, [0] * D
        for j, c in enumerate(trans):
            cond_map[c] = both[c] = (ops.COL_COND << 16) | j
        if ut is not None:
            for j, c in enumerate(ident):
                shared_map[c] = both[c] = (ops.COL_SHARED << 16) | j
        mk = lambda cmap, cm, cb, sm, sb: ops.SplineSpec(D, K, cmap, cm, cb, Pc, wh, sm, sb, Ps, list_mode)
        self._specs = {
            "both": mk(both, cmodes, cbounds, smodes, sbounds),        # density direction: one launch
            "shared": mk(shared_map, [], [], smodes, sbounds),         # sampling direction, step 1
            "cond": mk(cond_map, cmodes, cbounds, [], []),             # sampling direction, step 2
        }
        if img_ut:
            self._specs["ident_rows"] = self._ident_rows_spec
        self._id_cols = self.identity_features.to(torch.int32)

    def _cols(self, device):
        if self._id_cols.device != device:
            self._id_cols = self._id_cols.to(device)
        return self._id_cols

    # ---- images: every pixel is a row of C features; conditioner output [B, Ct*P, H, W] = rows [B*H*W, Ct*P] ----
    def _net_rows(self, identity_nchw, context):
        net = self.transform_net
        if context is None and hasattr(net, "forward_rows") and net.kernel_route(identity_nchw):
            return net.forward_rows(identity_nchw)
        out = net(identity_nchw, context)
        B, CP, H, W = out.shape
        return out.permute(0, 2, 3, 1).reshape(B * H * W, CP)

    def _image(self, inputs, context, inverse):
        if self._specs is None:
            self._build_specs()
        B, C, H, W = inputs.shape
        rows = inputs.permute(0, 2, 3, 1).reshape(B * H * W, C).contiguous()
        ut = self.unconditional_transform
        idf = self.identity_features
        logabsdet = 0.0

        def unconditional(rows):
            if not getattr(ut, "_img", False):
                raise ValueError("image inputs need img_shape for the unconditional transform (coupling.py:207-216)")
            ident = ops.gather_cols(rows, self._cols(rows.device))
            yi, ld = ops.rqs_layer(ident, ut.pixel_rows(B), None, self._specs["ident_rows"], inverse)
            rows = rows.index_copy(1, idf, yi)
            return rows, ld

        def net_params(rows):
            ident = rows.index_select(1, idf).reshape(B, H, W, len(idf)).permute(0, 3, 1, 2).contiguous()
            return self._net_rows(ident, context)
        if not inverse:
            params = net_params(rows)                                # conditioner sees the ORIGINAL identity part
            out, ld = ops.rqs_layer(rows, params, None, self._specs["cond"], False)
            if ut is not None:
                out, ld_i = unconditional(out)
                ld = ld + ld_i
        else:
            ld = 0.0
            if ut is not None:
                rows, ld = unconditional(rows)                       # coupling.py:108-110: identity part first
            params = net_params(rows)
            out, ld_c = ops.rqs_layer(rows, params, None, self._specs["cond"], True)
            ld = ld_c + ld
        return out.reshape(B, H, W, C).permute(0, 3, 1, 2).contiguous(), ld.reshape(B, H * W).sum(1) + logabsdet

    def forward(self, inputs, context=None, want_bins=False):
        self._check(inputs)
        self._check_domain(inputs)
        if inputs.dim() == 4:
            return self._image(inputs, context, False)
        if self._specs is None:
            self._build_specs()
        fused = None if want_bins else self._fused_route(inputs, context)
        if fused is not None and not self._needs_grad(inputs):
            return ops.resnet_spline(inputs, fused, *self.transform_net.fused_params(), inverse=False,
                                     shared=self._shared_params())
        if fused is not None and self.transform_net.hidden_features <= ops.KSLICE:
            # training: values from the one fused launch, graph (and backward kernels) of the layer-by-layer path
            shared = self._shared_params()
            y, ld, params, acts, xid = ops.resnet_spline(
                inputs.detach(), fused, *self.transform_net.fused_params(), inverse=False, shared=shared,
                want_params="padded" if self.transform_net.padded_out else True, want_acts=True,
                want_xid="padded" if self._pad_rows else True)
            identity_split = ops.gather_cols(inputs, self._cols(inputs.device), self._pad_rows, pre=xid)
            params = self.transform_net(identity_split, None, pre=(acts, params))
            return ops.rqs_layer(inputs, params, shared, self._specs["both"], False, pre=(y, ld))
        identity_split = ops.gather_cols(inputs, self._cols(inputs.device), self._pad_rows)
        params = self.transform_net(identity_split, context)
        shared = self.unconditional_transform.packed_params() if self.unconditional_transform is not None else None
        return ops.rqs_layer(inputs, params, shared, self._specs["both"], False, want_bins)

    def inverse(self, inputs, context=None):
        self._check(inputs)
        self._check_domain(inputs)
        if inputs.dim() == 4:
            return self._image(inputs, context, True)
        if self._specs is None:
            self._build_specs()
        fused = self._fused_route(inputs, context)
        if fused is not None and not self._needs_grad(inputs):
            return ops.resnet_spline(inputs, fused, *self.transform_net.fused_params(), inverse=True,
                                     shared=self._shared_params())
        logabsdet = 0.0
        if self.unconditional_transform is not None:
            inputs, logabsdet = ops.rqs_layer(inputs, None, self.unconditional_transform.packed_params(),
                                              self._specs["shared"], True)
        identity_split = ops.gather_cols(inputs, self._cols(inputs.device), self._pad_rows)
        params = self.transform_net(identity_split, context)
        outputs, logabsdet_split = ops.rqs_layer(inputs, params, None, self._specs["cond"], True)
        return outputs, logabsdet + logabsdet_split
