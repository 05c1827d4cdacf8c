"""MLP conditioner (reference: normflow/nets/mlp.py:7-58).  Round-1 status: the dense layers
are library GEMMs (cuBLAS through nn.Linear); the fused in-kernel conditioner for the tiny nets
of cfg 1 / cfg 4 and the tcgen05 path for hidden >= 128 are the next kernels."""
from torch import nn

from .. import utils


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
        return self.net(x)
