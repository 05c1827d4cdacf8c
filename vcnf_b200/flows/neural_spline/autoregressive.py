"""Masked autoregressive rational-quadratic spline
(reference: normflow/flows/neural_spline/autoregressive.py:18-137)."""
import numpy as np
import torch
from torch.nn import functional as F

from ..affine.autoregressive import Autoregressive
from .coupling import _modes_and_bounds, _params_per_feature, check_unit_box
from ... import ops
from ...nets import made as made_module
from ...utils import splines
from ...utils.nn import PeriodicFeatures


class MaskedPiecewiseRationalQuadraticAutoregressive(Autoregressive):
    """MADE conditioner + spline on every feature; the elementwise part (`_elementwise`,
    autoregressive.py:92-131) is one `vcnf_rqs_layer_fwd` launch with all D columns conditional.
    The MADE output is NOT divided by sqrt(hidden) (MADE has no `hidden_features`, hazard A.8)."""

    def __init__(self, features, hidden_features, context_features=None, num_bins=10, tails=None,
                 tail_bound=1., num_blocks=2, use_residual_blocks=True, random_mask=False,
                 permute_mask=False, activation=F.relu, dropout_probability=0., use_batch_norm=False,
                 init_identity=True, min_bin_width=splines.DEFAULT_MIN_BIN_WIDTH,
                 min_bin_height=splines.DEFAULT_MIN_BIN_HEIGHT,
                 min_derivative=splines.DEFAULT_MIN_DERIVATIVE):
        splines.check_minimums(num_bins, min_bin_width, min_bin_height, min_derivative)
        self.num_bins = num_bins
        self.tails = tails
        self.features = features
        if isinstance(self.tails, (list, tuple)):
            ind_circ = [i for i in range(features) if self.tails[i] == "circular"]
            if torch.is_tensor(tail_bound):
                scale_pf = np.pi / tail_bound[ind_circ]
            else:
                scale_pf = np.pi / tail_bound
            preprocessing = PeriodicFeatures(features, ind_circ, scale_pf)
        else:
            preprocessing = None
        autoregressive_net = made_module.MADE(
            features=features, hidden_features=hidden_features, context_features=context_features,
            num_blocks=num_blocks, output_multiplier=self._output_dim_multiplier(),
            use_residual_blocks=use_residual_blocks, random_mask=random_mask, permute_mask=permute_mask,
            activation=activation, dropout_probability=dropout_probability,
            use_batch_norm=use_batch_norm, preprocessing=preprocessing)
        if init_identity:
            torch.nn.init.constant_(autoregressive_net.final_layer.weight, 0.)
            torch.nn.init.constant_(autoregressive_net.final_layer.bias,
                                    np.log(np.exp(1 - min_derivative) - 1))
        super().__init__(autoregressive_net)
        if torch.is_tensor(tail_bound):
            self.register_buffer("tail_bound", tail_bound)
        else:
            self.tail_bound = tail_bound
        self._spec = None


    def _load_from_state_dict(self, *args, **kwargs):
        # the kernel specs cache the tail bounds: a loaded `tail_bound` buffer must rebuild them
        super()._load_from_state_dict(*args, **kwargs)
        self._spec = None
        self.__dict__.pop("_fused_cache", None)
        self.__dict__.pop("_feature_specs", None)
    def _output_dim_multiplier(self):
        return _params_per_feature(self.tails, self.num_bins)

    def forward(self, inputs, context=None):
        if self.tails is None:
            check_unit_box(inputs)                               # bounded spline: [0, 1] only (utils/splines.py:105-106)
        return super().forward(inputs, context)

    def inverse(self, inputs, context=None):
        if self.tails is None:
            check_unit_box(inputs)
        return super().inverse(inputs, context)

    def _elementwise(self, inputs, autoregressive_params, inverse=False):
        if self._spec is None:
            D, P = self.features, self._output_dim_multiplier()
            modes, bounds = _modes_and_bounds(self.tails, self.tail_bound, D)
            scale = 1.0
            if hasattr(self.autoregressive_net, "hidden_features"):
                scale = 1.0 / np.sqrt(self.autoregressive_net.hidden_features)
            cmap = [(ops.COL_COND << 16) | j for j in range(D)]
            self._spec = ops.SplineSpec(D, self.num_bins, cmap, modes, bounds, P, scale, [], [], P,
                                        isinstance(self.tails, (list, tuple)))
        return ops.rqs_layer(inputs, autoregressive_params, None, self._spec, inverse)

    def _elementwise_forward(self, inputs, autoregressive_params):
        return self._elementwise(inputs, autoregressive_params)

    def _elementwise_inverse(self, inputs, autoregressive_params):
        return self._elementwise(inputs, autoregressive_params, inverse=True)

    def _fused_spec(self):
        if hasattr(self.autoregressive_net, "hidden_features"):
            return None                                      # (MADE has none: the parameters are not rescaled, A.8)
        D = self.features
        cache = self.__dict__.setdefault("_fused_cache", {})
        dev = self.autoregressive_net.initial_layer.weight.device
        if cache.get("dev") != dev:
            modes, bounds = _modes_and_bounds(self.tails, self.tail_bound, D)
            cache.update(dev=dev, modes=torch.tensor(modes, dtype=torch.uint8, device=dev),
                         bounds=torch.tensor(bounds, dtype=torch.float32, device=dev))
        return 1, self.num_bins, cache["modes"], cache["bounds"], isinstance(self.tails, (list, tuple))

    def _inverse_feature(self, x_f, params_f, f):
        """One column of `_elementwise_inverse` (affine/autoregressive.py `_inverse_sliced`): x_f [B, 1],
        params_f [B, P] = the parameters of feature f."""
        D, P = self.features, self._output_dim_multiplier()
        cache = self.__dict__.setdefault("_feature_specs", {})
        if "mb" not in cache:                                   # once: a tensor tail_bound is read back here
            cache["mb"] = _modes_and_bounds(self.tails, self.tail_bound, D)
        modes, bounds = cache["mb"]
        key = (int(modes[f]), float(bounds[f]))
        if key not in cache:
            scale = 1.0
            if hasattr(self.autoregressive_net, "hidden_features"):
                scale = 1.0 / np.sqrt(self.autoregressive_net.hidden_features)
            cache[key] = ops.SplineSpec(1, self.num_bins, [(ops.COL_COND << 16) | 0], [modes[f]], [bounds[f]], P, scale,
                                        [], [], P, isinstance(self.tails, (list, tuple)))
        return ops.rqs_layer(x_f, params_f, None, cache[key], True)[0]
