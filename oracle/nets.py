"""Oracle conditioners: MLP, ResidualNet, MADE, ConvNet2d, PeriodicFeatures (plain torch).

Module trees mirror the reference so `load_state_dict` works key for key.
"""
import torch
from torch import nn
import torch.nn.functional as F


class MLP(nn.Module):
    """`normflow/nets/mlp.py:7-58`: Linear + LeakyReLU(leaky) ... Linear [+ output_fn]."""

    def __init__(self, layers, leaky=0.0, score_scale=None, output_fn=None,
                 output_scale=None, init_zeros=False, dropout=None):
        super().__init__()
        mods = []
        for a, b in zip(layers[:-2], layers[1:-1]):
            mods += [nn.Linear(a, b), nn.LeakyReLU(leaky)]
        if dropout is not None:
            mods.append(nn.Dropout(p=dropout))
        last = nn.Linear(layers[-2], layers[-1])
        if init_zeros:                                       # mlp.py:37-39
            nn.init.zeros_(last.weight)
            nn.init.zeros_(last.bias)
        mods.append(last)
        if output_fn is not None:                            # mlp.py:40-55
            if score_scale is not None:
                mods.append(_ConstScale(score_scale))
            fn = {"sigmoid": nn.Sigmoid, "relu": nn.ReLU, "tanh": nn.Tanh}.get(output_fn)
            if fn is not None:
                mods.append(fn())
            if output_scale is not None:
                mods.append(_ConstScale(output_scale))
        self.net = nn.Sequential(*mods)

    def forward(self, x):
        return self.net(x)


class _ConstScale(nn.Module):
    def __init__(self, scale):
        super().__init__()
        self.scale = scale

    def forward(self, x):
        return x * self.scale


class PeriodicFeatures(nn.Module):
    """`normflow/utils/nn.py:59-119`: w0*sin(scale*x)+w1*cos(scale*x) on selected columns."""

    def __init__(self, ndim, ind, scale=1.0):
        super().__init__()
        ind = torch.as_tensor(ind, dtype=torch.long)
        rest = torch.tensor([i for i in range(ndim) if i not in ind.tolist()], dtype=torch.long)
        self.register_buffer("ind", ind)
        self.register_buffer("ind_", rest)
        perm = torch.cat((ind, rest))
        self.register_buffer("inv_perm", torch.argsort(perm))
        self.weights = nn.Parameter(torch.ones(len(ind), 2))
        if torch.is_tensor(scale):
            self.register_buffer("scale", scale)
        else:
            self.scale = scale

    def forward(self, x):
        a = self.scale * x[..., self.ind]
        a = self.weights[:, 0] * torch.sin(a) + self.weights[:, 1] * torch.cos(a)
        return torch.cat((a, x[..., self.ind_]), -1)[..., self.inv_perm]


class _ResBlock(nn.Module):
    """`normflow/nets/resnet.py:8-58` without batch-norm/context (never used on the path)."""

    def __init__(self, features, activation):
        super().__init__()
        self.activation = activation
        self.linear_layers = nn.ModuleList([nn.Linear(features, features) for _ in range(2)])
        nn.init.uniform_(self.linear_layers[1].weight, -1e-3, 1e-3)   # resnet.py:34-36
        nn.init.uniform_(self.linear_layers[1].bias, -1e-3, 1e-3)

    def forward(self, x):
        t = self.linear_layers[0](self.activation(x))
        t = self.linear_layers[1](self.activation(t))
        return x + t


class ResidualNet(nn.Module):
    """`normflow/nets/resnet.py:60-106`."""

    def __init__(self, in_features, out_features, hidden_features, num_blocks=2,
                 activation=F.relu, preprocessing=None):
        super().__init__()
        self.hidden_features = hidden_features               # resnet.py:74 (read by a13)
        self.preprocessing = preprocessing
        self.initial_layer = nn.Linear(in_features, hidden_features)
        self.blocks = nn.ModuleList([_ResBlock(hidden_features, activation) for _ in range(num_blocks)])
        self.final_layer = nn.Linear(hidden_features, out_features)

    def forward(self, x, context=None):
        if self.preprocessing is not None:
            x = self.preprocessing(x)
        x = self.initial_layer(x)
        for blk in self.blocks:
            x = blk(x)
        return self.final_layer(x)


def _degrees_in(features):
    return torch.arange(1, features + 1)                     # made.py:15-17


class MaskedLinear(nn.Linear):
    """`normflow/nets/made.py:20-80`: y = x (W*mask)^T + b; buffers `mask`, `degrees`."""

    def __init__(self, in_degrees, out_features, autoregressive_features, is_output,
                 out_degrees_=None):
        super().__init__(len(in_degrees), out_features)
        if is_output:                                        # made.py:53-60
            base = _degrees_in(autoregressive_features) if out_degrees_ is None else out_degrees_
            mult = out_features // autoregressive_features
            deg = base.reshape(-1, 1).repeat(1, mult).reshape(-1)      # utils/nn.py:122-128 `tile`
            mask = (deg[:, None] > in_degrees).float()
        else:                                                # made.py:71-75
            hi = max(1, autoregressive_features - 1)
            lo = min(1, autoregressive_features - 1)
            deg = torch.arange(out_features) % hi + lo
            mask = (deg[:, None] >= in_degrees).float()
        self.register_buffer("mask", mask)
        self.register_buffer("degrees", deg)

    def forward(self, x):
        return F.linear(x, self.weight * self.mask, self.bias)


class _MaskedResBlock(nn.Module):
    """`normflow/nets/made.py:138-212`."""

    def __init__(self, in_degrees, autoregressive_features, activation):
        super().__init__()
        l0 = MaskedLinear(in_degrees, len(in_degrees), autoregressive_features, False)
        l1 = MaskedLinear(l0.degrees, len(in_degrees), autoregressive_features, False)
        self.linear_layers = nn.ModuleList([l0, l1])
        self.degrees = l1.degrees
        self.activation = activation
        nn.init.uniform_(l1.weight, a=-1e-3, b=1e-3)         # made.py:192-194
        nn.init.uniform_(l1.bias, a=-1e-3, b=1e-3)

    def forward(self, x):
        t = self.linear_layers[0](self.activation(x))
        t = self.linear_layers[1](self.activation(t))
        return x + t


class MADE(nn.Module):
    """`normflow/nets/made.py:215-300`, residual blocks, sequential degrees."""

    def __init__(self, features, hidden_features, num_blocks=2, output_multiplier=1,
                 permute_mask=False, activation=F.relu, preprocessing=None):
        super().__init__()
        self.preprocessing = preprocessing
        in_deg = _degrees_in(features)
        if permute_mask:
            in_deg = in_deg[torch.randperm(features)]
        self.initial_layer = MaskedLinear(in_deg, hidden_features, features, False)
        blocks, prev = [], self.initial_layer.degrees
        for _ in range(num_blocks):
            blocks.append(_MaskedResBlock(prev, features, activation))
            prev = blocks[-1].degrees
        self.blocks = nn.ModuleList(blocks)
        self.final_layer = MaskedLinear(prev, features * output_multiplier, features, True,
                                        out_degrees_=in_deg)

    def forward(self, x, context=None):
        if self.preprocessing is not None:
            x = self.preprocessing(x)
        x = self.initial_layer(x)
        for blk in self.blocks:
            x = blk(x)
        return self.final_layer(x)


class ConvNet2d(nn.Module):
    """`normflow/nets/cnn.py:7-47` (actnorm=False path): conv/LeakyReLU stack, zero-init last."""

    def __init__(self, channels, kernel_size, leaky=0.0, init_zeros=True):
        super().__init__()
        mods = []
        n = len(kernel_size)
        for i in range(n - 1):
            mods += [nn.Conv2d(channels[i], channels[i + 1], kernel_size[i],
                               padding=kernel_size[i] // 2), nn.LeakyReLU(leaky)]
        last = nn.Conv2d(channels[n - 1], channels[n], kernel_size[n - 1],
                         padding=kernel_size[n - 1] // 2)
        if init_zeros:
            nn.init.zeros_(last.weight)
            nn.init.zeros_(last.bias)
        mods.append(last)
        self.net = nn.Sequential(*mods)

    def forward(self, x):
        return self.net(x)


def __getattr__(name):
    # `nf.nets.VariationalDequantization` (normflow/nets/__init__.py) lives in oracle.core
    if name in ("VariationalDequantization", "Dequantization"):
        from . import core
        return getattr(core, name)
    raise AttributeError(name)
