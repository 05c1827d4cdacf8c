"""MLP conditioner (reference: normflow/nets/mlp.py:7-58).  Nets whose widths are all <= 32 (the [1,32,32,2]
net of the README Real NVP, the [15,30,15] / [1,4,1] nets of the dequantisation flows; fp32 and fp64, any
LeakyReLU slope >= 0) are evaluated in registers by ONE kernel per direction (`ops.mlp_small`).  Wider fp32
nets with `leaky=0.0` run every dense layer on the tcgen05 3xTF32 kernel (`ops.linear`), the ReLU fused into
the load of the following layer; anything else keeps nn.Linear (library GEMM).  Same module tree /
state_dict keys (`net.0.weight`, ...) as the reference either way."""
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

    def _fused_prefix(self):
        """Leading `Linear (LeakyReLU Linear)*` run of self.net with one common slope, and how many modules it spans."""
        mods = list(self.net)
        lin, slope, k = [mods[0]], None, 1
        while k + 1 < len(mods) and isinstance(mods[k], nn.LeakyReLU) and isinstance(mods[k + 1], nn.Linear):
            if slope is not None and mods[k].negative_slope != slope:
                break
            slope = mods[k].negative_slope
            lin.append(mods[k + 1])
            k += 2
        return lin, (0.0 if slope is None else slope), k

    def forward(self, x):
        if not (x.is_cuda and x.dim() == 2 and x.dtype in (torch.float32, torch.float64)):
            return self.net(x)
        mods = list(self.net)
        if isinstance(mods[0], nn.Linear):
            lin, slope, k = self._fused_prefix()
            weights, biases = [m.weight for m in lin], [m.bias for m in lin]
            if ops.mlp_small_supported(x, weights, slope):
                # the whole net in registers: one launch forward, one (+ reduce) backward -- vcnf_mlp_fwd / _bwd
                h = ops.mlp_small(x, weights, biases, slope)
                for m in mods[k:]:
                    h = m(h)
                return h
        if x.dtype != torch.float32:
            return self.net(x)
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
