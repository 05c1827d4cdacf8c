"""vcnf_b200 -- the normalizing-flow bijector hot path of telegraphroad/VCNF (a `normflow` fork)
as hand-written sm_100a CUDA kernels behind the reference's Python API.

    import vcnf_b200 as nf          # drop-in for `import normflow as nf` on the hot path
    model = nf.NormalizingFlow(nf.distributions.DiagGaussian(29), [nf.flows.CoupledRationalQuadraticSpline(29, 2, 128), ...])

The kernels live in libvcnf_b200.so (C ABI: include/vcnf_b200.h); there is no CPU fallback.
"""
from .core import NormalizingFlow, ClassCondFlow, MultiscaleFlow
from . import flows
from . import distributions
from . import transforms
from . import nets
from . import utils
from . import ops

__version__ = "0.1"
from . import parallel
from . import optim
