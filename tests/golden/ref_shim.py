"""Import the UNMODIFIED reference (`/root/reference/normflow`) in this container.

Test infrastructure only (golden-vector generation).  The reference
imports three modules that are absent here (`eagerpy`, `torch._six`, `pyro`); none of
them is touched by the bijector hot path, so inert stand-ins are registered before
the import.  `/root/reference` does not exist on the GPU box: nothing under `-m gpu`,
`smoke()` or `bench.py` imports this file.
"""
import math
import sys
import types

import torch

REFERENCE_ROOT = "/root/reference"


def load_reference():
    if "normflow" in sys.modules:
        return sys.modules["normflow"]
    eagerpy = types.ModuleType("eagerpy")
    eagerpy.squeeze = lambda *a, **k: None            # normflow/core.py:1 (never called)
    sys.modules.setdefault("eagerpy", eagerpy)
    six = types.ModuleType("torch._six")
    six.inf, six.nan = math.inf, math.nan             # distributions/base.py:16
    sys.modules.setdefault("torch._six", six)
    pyro = types.ModuleType("pyro")
    pdist = types.ModuleType("pyro.distributions")
    pdist.Chi2 = torch.distributions.Chi2             # distributions/base.py:15,24
    pdist.TorchDistribution = torch.distributions.Distribution
    pdist.MultivariateStudentT = torch.distributions.Distribution
    pyro.distributions = pdist
    sys.modules.setdefault("pyro", pyro)
    sys.modules.setdefault("pyro.distributions", pdist)
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    import normflow
    return normflow
