"""Reshaping layers (reference: normflow/flows/reshape.py)."""
import torch

from .base import Flow
from .. import ops


class Split(Flow):
    """Channel split (reshape.py:9-55).  'channel': z1 = first ceil(C/2) channels; 'channel_inv'
    flips the roles.  Inside AffineCouplingBlock the split never materialises (the coupling kernel
    indexes the halves in place); stand-alone (MultiscaleFlow levels) it is a view / one cat.
    'checkerboard' / 'checkerboard_inv' (reshape.py:30-44,56-72): z1 takes the positions whose non-batch indices
    sum to an odd number (even for '_inv'), z2 the rest, last dimension halved -- one index-map kernel each way
    (`vcnf_checkerboard`), no mask tensors, no `nonzero` synchronisation."""

    def __init__(self, mode="channel"):
        super().__init__()
        self.mode = mode

    def _check(self):
        if self.mode not in ("channel", "channel_inv") and "checkerboard" not in self.mode:
            raise NotImplementedError("Mode " + self.mode + " is not implemented.")

    def forward(self, z):
        self._check()
        if "checkerboard" in self.mode:
            z1, z2 = ops.checkerboard_split(z, "inv" in self.mode)
            return [z1, z2], 0
        a, b = z.chunk(2, dim=1)
        return ([a, b] if self.mode == "channel" else [b, a]), 0

    def inverse(self, z):
        self._check()
        z1, z2 = z
        if "checkerboard" in self.mode:
            return ops.checkerboard_merge(z1, z2, "inv" in self.mode), 0
        return torch.cat([z1, z2] if self.mode == "channel" else [z2, z1], 1), 0


class Merge(Split):
    """Split with forward and inverse interchanged (reshape.py:75-90)."""

    def forward(self, z):
        return Split.inverse(self, z)

    def inverse(self, z):
        return Split.forward(self, z)


class Squeeze(Flow):
    """2x2 space <-> depth (reshape.py:93-117) as one index-map kernel (`vcnf_squeeze`)."""

    def forward(self, z):
        return ops.squeeze(z, 0), 0

    def inverse(self, z):
        return ops.squeeze(z, 1), 0
