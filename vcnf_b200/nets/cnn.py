"""Glow's convolutional conditioner (reference: normflow/nets/cnn.py:7-47): conv -> LeakyReLU -> ... -> conv
(zero-init), stride 1, 'same' padding, as flows/affine/glow.py:24-34 builds it with kernel sizes (3, 1, 3).

Two routes, both hand-written kernels (no cuDNN on either):

* ``precision = "fp32"`` (default; what the golden-vector parity tests run): every convolution is
  vcnf_im2col (LeakyReLU of the previous layer fused into the read) -> vcnf_linear_tf32x3, the fp32-faithful
  3xTF32 tcgen05 GEMM of the dense conditioners, with their hand-written backward (vcnf_col2im,
  vcnf_linear_wgrad_tf32x3).
* ``precision = "bf16"`` (BASELINE cfg 5: "bf16 conditioner on tensor cores"): ops.convnet3_bf16 -- implicit-GEMM
  convolutions on tcgen05 with bf16 operands and fp32 accumulation (vcnf_conv_gemm_bf16 / vcnf_conv_wgrad_bf16).
  Stated tolerance against the fp32 reference: 2e-2 of the output scale (bf16 has 8 mantissa bits).

Select with ``vcnf_b200.nets.cnn.set_precision("bf16")`` or ``VCNF_CONV_PRECISION=bf16``.  The nn.Conv2d modules
only hold the parameters (state_dict keys net.{0,2,4}.{weight,bias} as in the reference); they are never called on
fp32 CUDA inputs.  fp64 models keep the library convolution (no fp64 tensor-core path)."""
import os

import torch
from torch import nn
import torch.nn.functional as F

from .. import ops

_PRECISION = os.environ.get("VCNF_CONV_PRECISION", "fp32")


def set_precision(p):
    global _PRECISION
    if p not in ("fp32", "bf16"):
        raise ValueError("precision must be 'fp32' or 'bf16'")
    _PRECISION = p


def get_precision():
    return _PRECISION


class ConvNet2d(nn.Module):
    def __init__(self, channels, kernel_size, leaky=0.0, init_zeros=True, actnorm=False, weight_std=None):
        super().__init__()
        if actnorm:
            raise NotImplementedError("net_actnorm=True is broken in the reference (utils/nn.py:36, hazard A.15)")
        stack = []
        n = len(kernel_size)
        for i in range(n - 1):
            conv = nn.Conv2d(channels[i], channels[i + 1], kernel_size[i], padding=kernel_size[i] // 2, bias=True)
            if weight_std is not None:
                conv.weight.data.normal_(mean=0.0, std=weight_std)
            stack += [conv, nn.LeakyReLU(leaky)]
        head = nn.Conv2d(channels[n - 1], channels[n], kernel_size[n - 1], padding=kernel_size[n - 1] // 2)
        if init_zeros:
            nn.init.zeros_(head.weight)
            nn.init.zeros_(head.bias)
        stack.append(head)
        self.net = nn.Sequential(*stack)
        self.leaky = float(leaky)
        self.kernel_size = tuple(kernel_size)

    def _convs(self):
        return [m for m in self.net if isinstance(m, nn.Conv2d)]

    def _bf16_ok(self, x):
        convs = self._convs()
        return (self.kernel_size == (3, 1, 3) and len(convs) == 3 and convs[0].out_channels % 64 == 0
                and convs[0].out_channels == convs[1].out_channels and convs[0].out_channels <= 256
                and all(c.bias is not None for c in convs) and self.leaky >= 0.0
                and 128 % x.shape[3] == 0)

    def forward(self, x):
        convs = self._convs()
        if not x.is_cuda:
            raise RuntimeError("vcnf_b200: ConvNet2d has no CPU path (got a tensor on %s)" % x.device)
        if x.dtype != torch.float32 or convs[0].weight.dtype != torch.float32:
            return self.net(x)                     # fp64 models: library convolution (documented)
        if any(k not in (1, 3) for k in self.kernel_size):
            raise NotImplementedError("ConvNet2d kernels are written for kernel sizes 1 and 3")
        if _PRECISION == "bf16":
            if not self._bf16_ok(x):
                raise NotImplementedError("bf16 ConvNet2d needs kernel sizes (3, 1, 3), hidden channels % 64 == 0 and "
                                          "<= 256, image width dividing 128")
            c = convs
            return ops.convnet3_bf16(x, c[0].weight, c[0].bias, c[1].weight, c[1].bias, c[2].weight, c[2].bias,
                                     self.leaky)
        return self._forward_fp32(x, convs)

    def _forward_fp32(self, x, convs):
        B, _, H, W = x.shape
        cur, layout = x, "nchw"
        for i, conv in enumerate(convs):
            k, cin, cout = self.kernel_size[i], conv.in_channels, conv.out_channels
            act = i > 0
            wr = conv.weight.permute(0, 2, 3, 1).reshape(cout, k * k * cin)
            if k == 1 and layout == "nhwc" and (not act or self.leaky == 0.0):
                cur = ops.linear(cur, wr, conv.bias, relu_in=act, pad_out=cout % 4 != 0)
            else:
                kk = k * k * cin
                kp = kk if kk <= ops.KSLICE else (kk + 3) // 4 * 4
                col = ops.im2col(cur, layout, B, cin, H, W, k, kp, act=act, slope=self.leaky)
                if kp != kk:
                    wr = F.pad(wr, (0, kp - kk))
                cur = ops.linear(col, wr, conv.bias, pad_out=cout % 4 != 0)
            layout = "nhwc"
        cout = convs[-1].out_channels
        return cur.reshape(B, H, W, cout).permute(0, 3, 1, 2).contiguous()
