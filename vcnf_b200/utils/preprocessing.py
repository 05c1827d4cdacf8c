"""Data-loader transforms (reference: normflow/utils/preprocessing.py): Logit, Jitter, Scale.  Device tensors go
through one kernel (vcnf_logit_fwd without log-det, vcnf_axpb); host tensors (inside DataLoader workers) stay on the
host with the same arithmetic."""
import torch

from .. import ops


def _kernel_ok(x):
    return torch.is_tensor(x) and x.is_cuda and x.dtype in (torch.float32, torch.float64) and x.numel() > 0


class Logit():
    """logit(alpha + (1 - alpha) x) (preprocessing.py:5-22)."""

    def __init__(self, alpha=0):
        self.alpha = alpha

    def __call__(self, x):
        if _kernel_ok(x) and x.dim() >= 1:
            return ops.logit_transform(x.reshape(1, -1), self.alpha, 1 - self.alpha, True, False).reshape(x.shape)
        x_ = self.alpha + (1 - self.alpha) * x
        return torch.log(x_ / (1 - x_))

    def inverse(self, x):
        if _kernel_ok(x) and x.dim() >= 1:
            return ops.logit_transform(x.reshape(1, -1), self.alpha, 1 - self.alpha, False, False).reshape(x.shape)
        return (torch.sigmoid(x) - self.alpha) / (1 - self.alpha)


class Jitter():
    """x + scale * U(0, 1) (preprocessing.py:25-39)."""

    def __init__(self, scale=1. / 256):
        self.scale = scale

    def __call__(self, x):
        eps = torch.rand_like(x)
        if _kernel_ok(x):
            return ops.axpb(x, 1.0, 0.0, noise=eps, c=self.scale)
        return x + eps * self.scale


class Scale():
    """x * scale (preprocessing.py:42-56)."""

    def __init__(self, scale=255. / 256.):
        self.scale = scale

    def __call__(self, x):
        if _kernel_ok(x):
            return ops.axpb(x, self.scale)
        return x * self.scale
