"""Residual MLP conditioner of the coupling spline layers (reference: normflow/nets/resnet.py:8-106).
`hidden_features` is read by the coupling layer for the 1/sqrt(hidden) scaling (coupling.py:314)."""
import torch
from torch import nn
from torch.nn import functional as F, init

from .. import ops


class ResidualBlock(nn.Module):
    """x + W1 act(W0 act(x)) (pre-activation; resnet.py:38-58)."""

    def __init__(self, features, context_features, activation=F.relu, dropout_probability=0.,
                 use_batch_norm=False, zero_initialization=True):
        super().__init__()
        self.activation = activation
        self.use_batch_norm = use_batch_norm
        if use_batch_norm:
            self.batch_norm_layers = nn.ModuleList([nn.BatchNorm1d(features, eps=1e-3) for _ in range(2)])
        if context_features is not None:
            self.context_layer = nn.Linear(context_features, features)
        self.linear_layers = nn.ModuleList([nn.Linear(features, features) for _ in range(2)])
        self.dropout = nn.Dropout(p=dropout_probability)
        if zero_initialization:
            init.uniform_(self.linear_layers[-1].weight, -1e-3, 1e-3)
            init.uniform_(self.linear_layers[-1].bias, -1e-3, 1e-3)

    def _fast(self, inputs, context):
        """Both dense layers on the tcgen05 kernel with ReLU / bias / residual fused."""
        is_relu = self.activation is F.relu or self.activation is torch.relu or isinstance(self.activation, nn.ReLU)
        return (context is None and not self.use_batch_norm and is_relu
                and (self.dropout.p == 0. or not self.training)
                and ops.linear_supported(inputs, self.linear_layers[0].weight))

    def forward(self, inputs, context=None):
        if self._fast(inputs, context):
            l0, l1 = self.linear_layers
            if l0.bias is not None and l1.bias is not None and ops.resblock_supported(inputs, l0.weight):
                return ops.residual_block(inputs, l0.weight, l0.bias, l1.weight, l1.bias)
            t = ops.linear(inputs, l0.weight, l0.bias, relu_in=True)
            return ops.linear(t, l1.weight, l1.bias, residual=inputs, relu_in=True)
        h = inputs
        if self.use_batch_norm:
            h = self.batch_norm_layers[0](h)
        h = self.linear_layers[0](self.activation(h))
        if self.use_batch_norm:
            h = self.batch_norm_layers[1](h)
        h = self.linear_layers[1](self.dropout(self.activation(h)))
        if context is not None:
            h = F.glu(torch.cat((h, self.context_layer(context)), dim=1), dim=1)
        return inputs + h


class ResidualNet(nn.Module):
    def __init__(self, in_features, out_features, hidden_features, context_features=None, num_blocks=2,
                 activation=F.relu, dropout_probability=0., use_batch_norm=False, preprocessing=None):
        super().__init__()
        self.hidden_features = hidden_features
        self.context_features = context_features
        self.preprocessing = preprocessing
        n_in = in_features if context_features is None else in_features + context_features
        self.initial_layer = nn.Linear(n_in, hidden_features)
        self.blocks = nn.ModuleList([
            ResidualBlock(features=hidden_features, context_features=context_features, activation=activation,
                          dropout_probability=dropout_probability, use_batch_norm=use_batch_norm)
            for _ in range(num_blocks)])
        self.final_layer = nn.Linear(hidden_features, out_features)
        # Set by the spline coupling layer that owns this net: its kernels take a leading dimension, so
        # the (3K-1)*T wide output may be a view with a 16-byte-multiple row pitch (TMA-movable tiles).
        # Stand-alone the net returns a contiguous tensor like the reference's.
        self.padded_out = False

    accepts_padded_rows = True   # inputs may be [B, n] views of wider storage (unit column stride)

    def forward(self, inputs, context=None):
        h = inputs if self.preprocessing is None else self.preprocessing(inputs)
        h = h if context is None else torch.cat((h, context), dim=1)
        fast = (ops.linear_supported(h, self.initial_layer.weight)
                and (self.hidden_features <= ops.KSLICE or self.hidden_features % 4 == 0))
        h = ops.linear(h, self.initial_layer.weight, self.initial_layer.bias) if fast else self.initial_layer(h)
        for block in self.blocks:
            h = block(h, context=context)
        if fast:
            return ops.linear(h, self.final_layer.weight, self.final_layer.bias, pad_out=self.padded_out)
        return self.final_layer(h)
