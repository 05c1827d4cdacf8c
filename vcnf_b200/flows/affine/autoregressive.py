"""Autoregressive base + masked affine autoregressive flow
(reference: normflow/flows/affine/autoregressive.py:13-103)."""
import numpy as np
import torch
from torch.nn import functional as F

from ..base import Flow
from ... import ops
from ...nets import made as made_module


class Autoregressive(Flow):
    """forward = one conditioner pass + elementwise kernel; inverse = D sequential passes, each
    feeding the next conditioner evaluation (autoregressive.py:24-36).  All passes stay in the
    autograd graph, exactly like the reference."""

    def __init__(self, autoregressive_net):
        super().__init__()
        self.autoregressive_net = autoregressive_net

    def forward(self, inputs, context=None):
        params = self.autoregressive_net(inputs, context)
        return self._elementwise_forward(inputs, params)

    # Sampling direction without a gradient.  The reference runs D full conditioner passes; pass t can only fix the
    # feature of degree t + 1, whose parameters are rows [f P, (f + 1) P) of the final masked layer and depend on
    # lower-degree features alone (made.py:56-60: strict mask at the output).  So passes 0 .. D-2 evaluate the hidden
    # layers, THOSE rows and ONE column of the elementwise inverse; the last pass is the reference's full pass and
    # yields every output and the log-determinant.  Same dot products, same values (a masked weight is an exact
    # zero), 1 / D of the final layer's work and of the [B, D P] parameter traffic in D - 1 of D passes.
    # `full_inverse = True` keeps the reference's loop (tests compare the two).
    full_inverse = False
    # One more exact reduction (default): the hidden units of degree t are final after pass t, so the whole D-pass
    # loop needs ONE evaluation of MADE, done band by band inside a single kernel launch (vcnf_made_inverse,
    # csrc/made_kernels.cu).  Same dot products in a different summation order: equal to the loop within 1e-5.
    # `fused_inverse = False` keeps the sliced passes above (bit-identical to the reference's loop).
    fused_inverse = True

    def _sliced_inverse_ok(self, inputs, context):
        net = self.autoregressive_net
        if self.full_inverse or context is not None or inputs.dim() != 2 or not hasattr(self, "_inverse_feature"):
            return False
        if not isinstance(net, made_module.MADE) or not hasattr(net, "final_rows"):
            return False
        if torch.is_grad_enabled() and (inputs.requires_grad or any(p.requires_grad for p in net.parameters())):
            return False
        for m in net.modules():
            if isinstance(m, torch.nn.modules.batchnorm._BatchNorm):
                return False
            if isinstance(m, torch.nn.Dropout) and m.p > 0. and m.training:
                return False
        return inputs.shape[1] > 1

    def _degree_order(self):
        """Features sorted by input degree (sequential, or permuted with `permute_mask`); cached host list."""
        deg = self.autoregressive_net.final_layer.degrees
        key = (deg.data_ptr(), deg._version)
        if getattr(self, "_order_key", None) != key:
            P = self._output_dim_multiplier()
            self._order = torch.argsort(deg[::P].cpu(), stable=True).tolist()
            self._order_key = key
        return self._order

    def _inverse_sliced(self, inputs):
        net, P = self.autoregressive_net, self._output_dim_multiplier()
        order = self._degree_order()
        outputs = torch.zeros_like(inputs)
        with torch.no_grad(), net.frozen_masks():
            for f in order[:-1]:
                h = net.hidden(outputs)
                params_f = net.final_rows(h, f * P, (f + 1) * P)
                outputs[:, f] = self._inverse_feature(inputs[:, f:f + 1], params_f, f)[:, 0]
            return self._elementwise_inverse(inputs, net(outputs))

    def _fused_spec(self):
        """(map code, K, modes, bounds, outside_zero) of the elementwise inverse for vcnf_made_inverse; None when the
        layer's map is not one the kernel evaluates."""
        return None

    def _inverse_fused(self, inputs):
        net = self.autoregressive_net
        spec = self._fused_spec()
        if spec is None or not self.fused_inverse or inputs.dtype != torch.float32 or not net.sampling_kernel_ok():
            return None
        packed = net.pack_for_sampling()
        if packed is None:
            return None
        map_code, K, modes, bounds, oz = spec
        return ops.made_inverse(inputs, packed, map_code, K, modes, bounds, oz)

    def inverse(self, inputs, context=None):
        if self._sliced_inverse_ok(inputs, context):
            out = self._inverse_fused(inputs)
            if out is not None:
                return out
            return self._inverse_sliced(inputs)
        num_inputs = int(np.prod(inputs.shape[1:]))
        outputs = torch.zeros_like(inputs)
        logabsdet = None
        for _ in range(num_inputs):
            params = self.autoregressive_net(outputs, context)
            outputs, logabsdet = self._elementwise_inverse(inputs, params)
        return outputs, logabsdet

    def _output_dim_multiplier(self):
        raise NotImplementedError()

    def _elementwise_forward(self, inputs, autoregressive_params):
        raise NotImplementedError()

    def _elementwise_inverse(self, inputs, autoregressive_params):
        raise NotImplementedError()


class MaskedAffineAutoregressive(Autoregressive):
    """MAF layer: scale = sigmoid(u+2)+1e-3, params viewed [B, D, 2] = (u, shift)
    (autoregressive.py:48-103); elementwise + log-det in `vcnf_affine_coupling_*` (MAF map)."""

    def __init__(self, features, hidden_features, context_features=None, num_blocks=2,
                 use_residual_blocks=True, random_mask=False, activation=F.relu,
                 dropout_probability=0., use_batch_norm=False):
        self.features = features
        made = made_module.MADE(
            features=features, hidden_features=hidden_features, context_features=context_features,
            num_blocks=num_blocks, output_multiplier=self._output_dim_multiplier(),
            use_residual_blocks=use_residual_blocks, random_mask=random_mask, activation=activation,
            dropout_probability=dropout_probability, use_batch_norm=use_batch_norm)
        super().__init__(made)

    def _output_dim_multiplier(self):
        return 2

    def _elementwise_forward(self, inputs, autoregressive_params):
        return ops.affine_coupling(inputs, autoregressive_params, 0, self.features,
                                   ops.SCALE_MAPS["maf"], False, swap_ts=True)

    def _elementwise_inverse(self, inputs, autoregressive_params):
        return ops.affine_coupling(inputs, autoregressive_params, 0, self.features,
                                   ops.SCALE_MAPS["maf"], True, swap_ts=True)

    def _fused_spec(self):
        return 0, 0, None, None, False

    def _inverse_feature(self, x_f, params_f, f):
        """One column of `_elementwise_inverse`: x_f [B, 1], params_f [B, 2] = (u, shift) of feature f."""
        return ops.affine_coupling(x_f.contiguous(), params_f, 0, 1, ops.SCALE_MAPS["maf"], True, swap_ts=True)[0]
