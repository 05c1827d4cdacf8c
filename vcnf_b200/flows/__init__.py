"""Bijector layers with the reference's public names (normflow/flows/__init__.py)."""
from .base import Flow, Reverse, Composite
from .reshape import Merge, Split, Squeeze
from .mixing import Permute, Invertible1x1Conv, LULinearPermute
from . import affine
from .affine.coupling import AffineConstFlow, AffineCoupling, MaskedAffineFlow, AffineCouplingBlock
from .affine.glow import GlowBlock
from .affine.autoregressive import MaskedAffineAutoregressive
from .normalization import ActNorm
from . import neural_spline
from .neural_spline import CoupledRationalQuadraticSpline, AutoregressiveRationalQuadraticSpline, \
    CircularCoupledRationalQuadraticSpline, CircularAutoregressiveRationalQuadraticSpline
