"""Residual conditioners of the coupling spline layers (reference: normflow/nets/resnet.py): the dense ResidualNet
(:8-106) and the convolutional ConvResidualNet of the image path (:109-213).  `hidden_features` / `hidden_channels`
are read by the coupling layer for the 1/sqrt(hidden) scaling (coupling.py:314-316)."""
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

    def forward(self, inputs, context=None, pre=None):
        """pre = (t, out) already computed by the fused net + spline kernel: only the autograd nodes are built."""
        if self._fast(inputs, context):
            l0, l1 = self.linear_layers
            if l0.bias is not None and l1.bias is not None and ops.resblock_supported(inputs, l0.weight):
                return ops.residual_block(inputs, l0.weight, l0.bias, l1.weight, l1.bias, pre=pre)
            t = ops.linear(inputs, l0.weight, l0.bias, relu_in=True, pre=None if pre is None else pre[0])
            return ops.linear(t, l1.weight, l1.bias, residual=inputs, relu_in=True, pre=None if pre is None else pre[1])
        if pre is not None:
            raise RuntimeError("ResidualBlock: precomputed outputs need the kernel route")
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

    def fusable(self):
        """True when the whole net fits vcnf_resnet_spline_fwd (ops.resnet_spline): plain ReLU residual blocks of width
        128 on fp32 CUDA parameters, nothing stochastic."""
        if self.context_features is not None or self.preprocessing is not None or self.hidden_features != 128:
            return False
        if not (0 <= len(self.blocks) <= 4):
            return False
        for blk in self.blocks:
            if blk.use_batch_norm or not _is_relu(blk.activation) or (blk.dropout.p != 0. and blk.training):
                return False
        w = self.initial_layer.weight
        return w.is_cuda and w.dtype == torch.float32

    def fused_params(self):
        layers = [self.initial_layer] + [l for blk in self.blocks for l in blk.linear_layers] + [self.final_layer]
        return ([l.weight.detach().contiguous() for l in layers],
                [None if l.bias is None else l.bias.detach().contiguous() for l in layers])

    def forward(self, inputs, context=None, pre=None):
        """pre = (acts [1 + 2 blocks, B, hidden], params) from ops.resnet_spline: the values are taken from there and only
        the autograd nodes of the layer-by-layer path are built (their backward kernels run as usual)."""
        h = inputs if self.preprocessing is None else self.preprocessing(inputs)
        h = h if context is None else torch.cat((h, context), dim=1)
        fast = (ops.linear_supported(h, self.initial_layer.weight)
                and (self.hidden_features <= ops.KSLICE or self.hidden_features % 4 == 0))
        if pre is not None:
            if not fast:
                raise RuntimeError("ResidualNet: precomputed outputs need the kernel route")
            acts, params = pre
            h = ops.linear(h, self.initial_layer.weight, self.initial_layer.bias, pre=acts[0])
            for i, block in enumerate(self.blocks):
                h = block(h, context=context, pre=(acts[2 * i + 1], acts[2 * i + 2]))
            return ops.linear(h, self.final_layer.weight, self.final_layer.bias, pad_out=self.padded_out, pre=params)
        h = ops.linear(h, self.initial_layer.weight, self.initial_layer.bias) if fast else self.initial_layer(h)
        for block in self.blocks:
            h = block(h, context=context)
        if fast:
            return ops.linear(h, self.final_layer.weight, self.final_layer.bias, pad_out=self.padded_out)
        return self.final_layer(h)


def _is_relu(fn):
    return fn is F.relu or fn is torch.relu or isinstance(fn, nn.ReLU)


class ConvResidualBlock(nn.Module):
    """x + conv3x3(act(conv3x3(act(x)))) (resnet.py:109-161).  On fp32 CUDA inputs each convolution is
    vcnf_im2col (the ReLU fused into the read) -> vcnf_linear_tf32x3 on NHWC pixel rows, the skip connection rides on
    the second GEMM's residual operand; no library convolution."""

    def __init__(self, channels, context_channels=None, activation=F.relu, dropout_probability=0.,
                 use_batch_norm=False, zero_initialization=True):
        super().__init__()
        self.activation = activation
        if context_channels is not None:
            self.context_layer = nn.Conv2d(context_channels, channels, kernel_size=1, padding=0)
        self.use_batch_norm = use_batch_norm
        if use_batch_norm:
            self.batch_norm_layers = nn.ModuleList([nn.BatchNorm2d(channels, eps=1e-3) for _ in range(2)])
        self.conv_layers = nn.ModuleList([nn.Conv2d(channels, channels, kernel_size=3, padding=1) for _ in range(2)])
        self.dropout = nn.Dropout(p=dropout_probability)
        if zero_initialization:
            init.uniform_(self.conv_layers[-1].weight, -1e-3, 1e-3)
            init.uniform_(self.conv_layers[-1].bias, -1e-3, 1e-3)

    def kernel_route(self, context):
        return (context is None and not self.use_batch_norm and _is_relu(self.activation)
                and (self.dropout.p == 0. or not self.training))

    def forward_rows(self, rows, B, H, W):
        """rows: NHWC pixel rows [B*H*W, C] fp32."""
        C = self.conv_layers[0].in_channels
        kk = 9 * C
        kp = kk if kk <= ops.KSLICE else (kk + 3) // 4 * 4
        cur = rows
        for i, conv in enumerate(self.conv_layers):
            wr = conv.weight.permute(0, 2, 3, 1).reshape(C, kk)
            if kp != kk:
                wr = F.pad(wr, (0, kp - kk))
            col = ops.im2col(cur, "nhwc", B, C, H, W, 3, kp, act=True, slope=0.0)
            cur = ops.linear(col, wr, conv.bias, residual=rows if i == 1 else None, pad_out=C % 4 != 0)
        return cur

    def forward(self, inputs, context=None):
        temps = inputs
        if self.use_batch_norm:
            temps = self.batch_norm_layers[0](temps)
        temps = self.conv_layers[0](self.activation(temps))
        if self.use_batch_norm:
            temps = self.batch_norm_layers[1](temps)
        temps = self.conv_layers[1](self.dropout(self.activation(temps)))
        if context is not None:
            temps = F.glu(torch.cat((temps, self.context_layer(context)), dim=1), dim=1)
        return inputs + temps


class ConvResidualNet(nn.Module):
    """1x1 conv -> residual blocks -> 1x1 conv (resnet.py:164-213), the conditioner of the 4-D spline coupling.
    fp32 CUDA inputs without context / batch norm / active dropout run entirely on the package's kernels in NHWC
    (`forward_rows` hands the coupling layer the [B*H*W, out] parameter rows it consumes directly); everything else
    (fp64, context, batch norm) evaluates the modules as the reference does."""

    def __init__(self, in_channels, out_channels, hidden_channels, context_channels=None, num_blocks=2,
                 activation=F.relu, dropout_probability=0., use_batch_norm=False):
        super().__init__()
        self.context_channels = context_channels
        self.hidden_channels = hidden_channels
        n_in = in_channels if context_channels is None else in_channels + context_channels
        self.initial_layer = nn.Conv2d(n_in, hidden_channels, kernel_size=1, padding=0)
        self.blocks = nn.ModuleList([
            ConvResidualBlock(channels=hidden_channels, context_channels=context_channels, activation=activation,
                              dropout_probability=dropout_probability, use_batch_norm=use_batch_norm)
            for _ in range(num_blocks)])
        self.final_layer = nn.Conv2d(hidden_channels, out_channels, kernel_size=1, padding=0)

    def kernel_route(self, inputs, context=None):
        return (inputs.is_cuda and inputs.dtype == torch.float32 and inputs.dim() == 4
                and self.initial_layer.weight.dtype == torch.float32
                and all(b.kernel_route(context) for b in self.blocks) and context is None)

    def forward_rows(self, inputs):
        """NCHW fp32 [B, C, H, W] -> parameter rows [B*H*W, out_channels] (NHWC)."""
        B, C, H, W = inputs.shape
        hid = self.hidden_channels
        rows = ops.im2col(inputs, "nchw", B, C, H, W, 1)                      # layout change only (1x1 patches)
        cur = ops.linear(rows, self.initial_layer.weight.reshape(hid, C), self.initial_layer.bias,
                         pad_out=hid % 4 != 0)
        for block in self.blocks:
            cur = block.forward_rows(cur, B, H, W)
        out_c = self.final_layer.out_channels
        return ops.linear(cur, self.final_layer.weight.reshape(out_c, hid), self.final_layer.bias,
                          pad_out=out_c % 4 != 0)

    def forward(self, inputs, context=None):
        if self.kernel_route(inputs, context):
            B, _, H, W = inputs.shape
            rows = self.forward_rows(inputs)
            return rows.reshape(B, H, W, -1).permute(0, 3, 1, 2).contiguous()
        temps = self.initial_layer(inputs if context is None else torch.cat((inputs, context), dim=1))
        for block in self.blocks:
            temps = block(temps, context)
        return self.final_layer(temps)
