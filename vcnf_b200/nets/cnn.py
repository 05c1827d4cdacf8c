"""Glow's convolutional conditioner (reference: normflow/nets/cnn.py:7-47): conv3x3 -> LeakyReLU
-> conv1x1 -> LeakyReLU -> conv3x3 (zero-init).  Round-1 status: library convolutions (cuDNN) in
full fp32 -- cuDNN's default TF32 convolutions are switched off for these calls because they miss the
stated 1e-5 / 1e-4 parity tolerance (measured: 79 x tolerance on the golden Glow block); the bf16
tcgen05 implicit-GEMM path is the next kernel for cfg 5."""
import torch
from torch import nn


class ConvNet2d(nn.Module):
    def __init__(self, channels, kernel_size, leaky=0.0, init_zeros=True, actnorm=False, weight_std=None):
        super().__init__()
        if actnorm:
            raise NotImplementedError("net_actnorm=True is broken in the reference (utils/nn.py:36, hazard A.15)")
        # fp32 parity mode for the library convolutions, forward AND backward (the backward runs
        # later under autograd, so a forward-only context manager would not cover it)
        torch.backends.cudnn.allow_tf32 = False
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

    def forward(self, x):
        return self.net(x)
