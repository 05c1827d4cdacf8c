"""vcnf_const_affine_bwd_static: the no-allocation variant of vcnf_const_affine_bwd (scratch in a file-scope
device array, include/vcnf_b200.h) must give the existing entry point's results, call after call.  Collected last
on purpose."""
import pytest
import torch

from vcnf_b200 import _lib

pytestmark = pytest.mark.gpu


def _run(name, z, s, t, g, S, inverse, ws=None):
    B, Pn = z.shape[0], s.numel()
    gz, gs, gt = torch.empty_like(z), torch.empty_like(s), torch.empty_like(t)
    args = [_lib.ptr(z), _lib.ptr(s), _lib.ptr(t), _lib.ptr(g), _lib.ptr(gz), _lib.ptr(gs), _lib.ptr(gt)]
    if ws is not None:
        args.append(_lib.ptr(ws))
    args += [B, Pn, S, int(inverse), _lib.dtype_code(z), _lib.stream()]
    _lib.call(name, *args)
    return gz, gs, gt


# the second call is smaller than the first and the third a different dtype: the shared array then holds stale
# partials of an earlier call beside the ones in use
CASES = [(4099, 29, 1, 0, torch.float32), (513, 15, 1, 1, torch.float32), (64, 12, 16, 1, torch.float64),
         (300, 256, 1, 0, torch.float32)]


def test_static_scratch_matches_workspace_entry_point_on_consecutive_calls():
    dev = torch.device("cuda:0")
    gen = torch.Generator().manual_seed(7)
    for rep in range(2):
        for (B, Pn, S, inverse, dt) in CASES:
            z = torch.randn(B, Pn, S, generator=gen, dtype=dt).to(dev)
            g = torch.randn(B, Pn, S, generator=gen, dtype=dt).to(dev)
            s = (0.3 * torch.randn(Pn, generator=gen, dtype=dt)).to(dev)
            t = torch.randn(Pn, generator=gen, dtype=dt).to(dev)
            ws = torch.empty(_lib.query("vcnf_const_affine_bwd_workspace", Pn, _lib.dtype_code(z)),
                             dtype=torch.uint8, device=dev)
            want = _run("vcnf_const_affine_bwd", z, s, t, g, S, inverse, ws)
            got1 = _run("vcnf_const_affine_bwd_static", z, s, t, g, S, inverse)
            got2 = _run("vcnf_const_affine_bwd_static", z, s, t, g, S, inverse)   # two calls in a row
            torch.cuda.synchronize()
            for w, a, b in zip(want, got1, got2):
                # same kernels, same fixed summation order: bit-identical (the tolerance of the deterministic-
                # reduction tests in this suite)
                assert torch.equal(w, a) and torch.equal(w, b), (B, Pn, S, inverse, dt, rep)
            # and against a plain torch evaluation at the affine tolerance of tests/parity.py
            e = torch.exp(-s if inverse else s).view(1, Pn, 1)
            if inverse:
                ref_gs = -(g * (z - t.view(1, Pn, 1)) * e).sum((0, 2))
            else:
                ref_gs = (g * z * e).sum((0, 2))
            scale = float(ref_gs.abs().max()) + 1e-12
            assert float((got2[1] - ref_gs).abs().max()) <= (1e-9 if dt == torch.float64 else 2e-4) * scale


def test_static_scratch_rejects_sizes_beyond_its_maximum():
    dev = torch.device("cuda:0")
    Pn = 257
    z = torch.zeros(4, Pn, 2, device=dev)
    s = torch.zeros(Pn, device=dev)
    with pytest.raises(RuntimeError, match="unsupported size"):
        _run("vcnf_const_affine_bwd_static", z, s, s, z, 2, 0)
