"""Flow API (reference: normflow/flows/base.py:6-70): `forward(z) -> (z', log_det)`,
`inverse(z) -> (z', log_det)`."""
import torch
from torch import nn


class Flow(nn.Module):
    """Generic bijector; log_det is a [B] tensor, a 0-dim tensor or python 0 (base.py:12-21)."""

    def forward(self, z):
        raise NotImplementedError("Forward pass has not been implemented.")

    def inverse(self, z):
        raise NotImplementedError("This flow has no algebraic inverse.")


class Reverse(Flow):
    """Swaps forward and inverse of a flow (base.py:24-40)."""

    def __init__(self, flow):
        super().__init__()
        self.flow = flow

    def forward(self, z):
        return self.flow.inverse(z)

    def inverse(self, z):
        return self.flow.forward(z)


class Composite(Flow):
    """Chains flows (base.py:43-70).  Unlike the reference, whose accumulator is a CPU tensor
    (base.py:59), the accumulator lives on the input's device."""

    def __init__(self, flows):
        super().__init__()
        self._flows = nn.ModuleList(flows)

    @staticmethod
    def _cascade(inputs, funcs):
        total = torch.zeros(inputs.shape[0], dtype=inputs.dtype, device=inputs.device)
        out = inputs
        for f in funcs:
            out, ld = f(out)
            total = total + ld
        return out, total

    def forward(self, inputs):
        return self._cascade(inputs, self._flows)

    def inverse(self, inputs):
        return self._cascade(inputs, (f.inverse for f in self._flows[::-1]))
