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

    def inverse(self, inputs, context=None):
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
