from .base import BaseDistribution, DiagGaussian, ClassCondDiagGaussian, GaussianMixture
from .target import Target, TwoMoons
