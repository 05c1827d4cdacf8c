from .nn import ClampExp, ConstScaleLayer, PeriodicFeatures, tile, sum_except_batch
from .optim import clear_grad, set_requires_grad
from . import masks, splines
from .preprocessing import Logit, Jitter, Scale
