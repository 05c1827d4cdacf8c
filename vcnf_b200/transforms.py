"""Data transforms applied around the flow (reference: normflow/transforms.py) -- `Logit` is the optional
`MultiscaleFlow.transform` of the Glow / RealNVP image recipes (example/glow.ipynb), `Shift` re-centres [0, 1] data.
Both run as one kernel per call on the device (vcnf_logit_fwd / _bwd with the per-sample log-det reduced in the same
pass; vcnf_axpb in place)."""
import numpy as np
import torch

from . import flows, ops


class Logit(flows.Flow):
    """logit(alpha + (1 - 2 alpha) x) (transforms.py:7-39); `forward` is the generation direction (sigmoid)."""

    def __init__(self, alpha=0.05):
        super().__init__()
        self.alpha = alpha

    def forward(self, z):
        beta = 1 - 2 * self.alpha
        if z.is_cuda:
            return ops.logit_transform(z, self.alpha, beta, to_logit=False)
        dims = list(range(1, z.dim()))
        ls = torch.sum(torch.nn.functional.logsigmoid(z), dim=dims)
        mls = torch.sum(torch.nn.functional.logsigmoid(-z), dim=dims)
        return (torch.sigmoid(z) - self.alpha) / beta, -np.log(beta) * np.prod([*z.shape[1:]]) + ls + mls

    def inverse(self, z):
        beta = 1 - 2 * self.alpha
        if z.is_cuda:
            return ops.logit_transform(z, self.alpha, beta, to_logit=True)
        z = self.alpha + beta * z
        logz, log1mz = torch.log(z), torch.log(1 - z)
        dims = list(range(1, z.dim()))
        return logz - log1mz, np.log(beta) * np.prod([*z.shape[1:]]) - torch.sum(logz, dim=dims) - torch.sum(log1mz, dim=dims)


class Shift(flows.Flow):
    """z - shift / z + shift IN PLACE like the reference (transforms.py:42-62, SURVEY.md 8b aliasing)."""

    def __init__(self, shift=-0.5):
        super().__init__()
        self.shift = shift

    def _apply(self, z, delta):
        if z.is_cuda and z.is_contiguous() and not (torch.is_grad_enabled() and z.requires_grad) \
                and z.dtype in (torch.float32, torch.float64):
            return ops.axpb(z, 1.0, delta, out=z)
        z += delta
        return z

    def forward(self, z):
        return self._apply(z, -self.shift), 0.

    def inverse(self, z):
        return self._apply(z, self.shift), 0.
