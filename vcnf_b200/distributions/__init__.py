from . import base, extra
from .base import BaseDistribution, DiagGaussian, ClassCondDiagGaussian, GaussianMixture
from .extra import MultivariateGaussian, T, GGD, GMM, GenNormal
from .target import Target, TwoMoons, NealsFunnel

# the run scripts address them as nf.distributions.base.<name> (runcred.py:92-117)
for _n in ("MultivariateGaussian", "T", "GGD", "GMM", "GenNormal"):
    setattr(base, _n, getattr(extra, _n))
del _n
