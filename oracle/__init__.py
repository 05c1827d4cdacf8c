"""CPU oracle for the VCNF bijector hot path -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

A plain-PyTorch (CPU, fp32/fp64) restatement of the reference algorithm for the path
`BASELINE.json: north_star` names: coupling / autoregressive bijectors, their
conditioners, the per-sample log-det reduction and the `NormalizingFlow` containers.
Every function cites the reference file:line it follows (paths relative to the
reference root, `normflow/...`).

Who may import this package: `tests/`, `__graft_entry__.smoke()` and the
`cpu_baseline` / `--impl reference` legs of `bench.py` -- as the checker or the timed
CPU baseline only.  Nothing under `vcnf_b200/` imports it; the product path has no
CPU fallback and fails loudly when `libvcnf_b200.so` is missing.

Parity pinning: the reference's own tests hold no golden vectors, so the
oracle is pinned against outputs of the unmodified reference run in the build
container: `tests/golden/make_golden.py` imports `/root/reference/normflow`, evaluates
every layer on seeded inputs and commits the results to `tests/golden/*.pt`;
`tests/test_oracle_golden.py` checks the oracle against them (state dicts are loaded
key-for-key, so the module trees here mirror the reference's `state_dict` layout).
"""
from . import splines, nets, flows, distributions, core  # noqa: F401
from .core import NormalizingFlow, MultiscaleFlow  # noqa: F401
