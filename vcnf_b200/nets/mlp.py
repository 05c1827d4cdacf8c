"""MLP conditioner (reference: normflow/nets/mlp.py:7-58).  fp32 CUDA operands with the default
`leaky=0.0` run every dense layer on the tcgen05 3xTF32 kernel (`ops.linear`), the ReLU fused into the
load of the following layer; fp64 models (`nfm.double()`) and leaky slopes != 0 keep nn.Linear (library
GEMM).  Same module tree / state_dict keys (`net.0.weight`, ...) as the reference either way."""
import torch
from torch import nn

from .. import ops, utils


class MLP(nn.Module):
    def __init__(self, layers, leaky=0.0, score_scale=None, output_fn=None,
                 output_scale=None, init_zeros=False, dropout=None):
        super().__init__()
        stack = []
        for n_in, n_out in zip(layers[:-2], layers[1:-1]):
            stack.append(nn.Linear(n_in, n_out))
            stack.append(nn.LeakyReLU(leaky))
        if dropout is not None:
            stack.append(nn.Dropout(p=dropout))
        head = nn.Linear(layers[-2], layers[-1])
        if init_zeros:
            nn.init.zeros_(head.weight)
            nn.init.zeros_(head.bias)
        stack.append(head)
        if output_fn is not None:
            if score_scale is not None:
                stack.append(utils.ConstScaleLayer(score_scale))
            fns = {"sigmoid": nn.Sigmoid, "relu": nn.ReLU, "tanh": nn.Tanh, "clampexp": utils.ClampExp}
            if output_fn in fns:
                stack.append(fns[output_fn]())
            if output_scale is not None:
                stack.append(utils.ConstScaleLayer(output_scale))
        self.net = nn.Sequential(*stack)

    def forward(self, x):
        if not (x.is_cuda and x.dtype == torch.float32 and x.dim() == 2):
            return self.net(x)
        mods = list(self.net)
        h, pending_relu = x, False
        for i, m in enumerate(mods):
            if isinstance(m, nn.Linear) and ops.linear_supported(h, m.weight):
                h = ops.linear(h, m.weight, m.bias, relu_in=pending_relu)
                pending_relu = False
                continue
            if pending_relu:
                h, pending_relu = torch.relu(h), False
            if (isinstance(m, nn.LeakyReLU) and m.negative_slope == 0.0 and i + 1 < len(mods)
                    and isinstance(mods[i + 1], nn.Linear)):
                pending_relu = True      # applied by the next layer's operand load
            else:
                h = m(h)
        return torch.relu(h) if pending_relu else h
