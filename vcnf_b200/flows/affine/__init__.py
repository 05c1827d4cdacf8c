from . import coupling, autoregressive
from .coupling import AffineConstFlow, AffineCoupling, MaskedAffineFlow, AffineCouplingBlock
from .autoregressive import Autoregressive, MaskedAffineAutoregressive
