"""Golden-vector case table shared by the generator (run against the unmodified reference
in the build container) and the parity tests (run against the oracle on CPU and against
the CUDA product on the GPU).  Every builder takes a namespace with the reference's public
layout (`nf.flows.X`, `nf.nets.X`, `nf.distributions.X`) -- the reference, `oracle` and
`vcnf_b200` all provide it, which is the drop-in contract of SURVEY.md 8(b).
"""
import torch
from torch import nn


def _g(seed):
    g = torch.Generator()
    g.manual_seed(seed)
    return g


def randomize(module, seed=7, scale=0.3):
    """Perturb every floating-point parameter so zero-initialised layers do real work."""
    g = _g(seed)
    with torch.no_grad():
        for _, p in sorted(module.named_parameters()):
            p.add_(scale * torch.randn(p.shape, generator=g, dtype=torch.float32).to(p.dtype))
    return module


def _mlp_pair(nf, d, h):
    return nf.nets.MLP([d, h, d], init_zeros=True), nf.nets.MLP([d, h, d], init_zeros=True)


def _alt_mask(d, flip=False):
    b = torch.tensor([1.0 if i % 2 == 0 else 0.0 for i in range(d)])
    return 1 - b if flip else b


# name -> dict(build, x, dtype, directions, tol)
LAYER_CASES = {
    "crqs_linear": dict(
        build=lambda nf: nf.flows.CoupledRationalQuadraticSpline(29, 2, 32),
        x=lambda: 1.5 * torch.randn(96, 29, generator=_g(1))),
    "crqs_linear_k10_rev": dict(
        build=lambda nf: nf.flows.CoupledRationalQuadraticSpline(6, 1, 16, num_bins=10, tail_bound=2.0,
                                                                 reverse_mask=True),
        x=lambda: 1.2 * torch.randn(64, 6, generator=_g(2))),
    "crqs_circular_str": dict(
        build=lambda nf: nf.flows.CoupledRationalQuadraticSpline(5, 1, 16, tails="circular", tail_bound=3.0),
        x=lambda: 1.5 * torch.randn(64, 5, generator=_g(3))),
    "ccrqs_mixed": dict(
        build=lambda nf: nf.flows.CircularCoupledRationalQuadraticSpline(
            6, 1, 16, ind_circ=[1, 2, 4], tail_bound=torch.tensor([3.0, 3.14159, 3.14159, 2.0, 3.14159, 4.0])),
        x=lambda: 1.5 * torch.randn(64, 6, generator=_g(4))),
    "arqs": dict(
        build=lambda nf: nf.flows.AutoregressiveRationalQuadraticSpline(8, 2, 32),
        x=lambda: 1.5 * torch.randn(48, 8, generator=_g(5)), scale=0.15),
    "carqs_mixed": dict(
        build=lambda nf: nf.flows.CircularAutoregressiveRationalQuadraticSpline(
            5, 1, 16, ind_circ=[0, 3], tail_bound=3.0, permute_mask=False),
        x=lambda: 1.5 * torch.randn(32, 5, generator=_g(6)), scale=0.15),
    "lu_permute": dict(
        build=lambda nf: nf.flows.LULinearPermute(29, identity_init=False),
        x=lambda: torch.randn(64, 29, generator=_g(7))),
    "affine_block_exp": dict(
        build=lambda nf: nf.flows.AffineCouplingBlock(nf.nets.MLP([1, 32, 32, 2], init_zeros=True)),
        x=lambda: torch.randn(128, 2, generator=_g(8))),
    "affine_block_odd_sigmoid": dict(
        build=lambda nf: nf.flows.AffineCouplingBlock(nf.nets.MLP([3, 16, 4]), scale_map="sigmoid"),
        x=lambda: torch.randn(64, 5, generator=_g(9))),
    "affine_block_noscale_inv": dict(
        build=lambda nf: nf.flows.AffineCouplingBlock(nf.nets.MLP([2, 16, 3]), scale=False,
                                                      split_mode="channel_inv"),
        x=lambda: torch.randn(64, 5, generator=_g(10)), nograd=True),
    "permute_swap_odd": dict(
        build=lambda nf: nf.flows.Permute(5, mode="swap"),
        x=lambda: torch.randn(16, 5, generator=_g(11))),
    "permute_shuffle": dict(
        build=lambda nf: nf.flows.Permute(7, mode="shuffle"),
        x=lambda: torch.randn(16, 7, generator=_g(12))),
    "masked_affine_f64": dict(
        build=lambda nf: nf.flows.MaskedAffineFlow(_alt_mask(15), *_mlp_pair(nf, 15, 30)),
        x=lambda: torch.randn(64, 15, generator=_g(13), dtype=torch.float64), dtype=torch.float64),
    "masked_affine_f32": dict(
        build=lambda nf: nf.flows.MaskedAffineFlow(_alt_mask(29, True), *_mlp_pair(nf, 29, 58)),
        x=lambda: torch.randn(64, 29, generator=_g(14))),
    "actnorm_2d": dict(
        build=lambda nf: nf.flows.ActNorm(15),
        x=lambda: 2.0 * torch.randn(64, 15, generator=_g(15)) + 0.5, randomize=False),
    "actnorm_4d": dict(
        build=lambda nf: nf.flows.ActNorm((6, 1, 1)),
        x=lambda: 2.0 * torch.randn(8, 6, 4, 4, generator=_g(16)) + 0.5, randomize=False),
    "maf": dict(
        build=lambda nf: nf.flows.MaskedAffineAutoregressive(8, 32),
        x=lambda: torch.randn(48, 8, generator=_g(17)), scale=0.15),
    "inv1x1_lu": dict(
        build=lambda nf: nf.flows.Invertible1x1Conv(12, use_lu=True),
        x=lambda: torch.randn(4, 12, 4, 4, generator=_g(18)), scale=0.05),
    "squeeze": dict(
        build=lambda nf: nf.flows.Squeeze(),
        x=lambda: torch.randn(3, 8, 4, 6, generator=_g(19))),
    "glow_block": dict(
        build=lambda nf: nf.flows.GlowBlock(12, 32),
        x=lambda: torch.randn(4, 12, 8, 8, generator=_g(20)), scale=0.05),
}


def _img_coupling(nf, uncond):
    cpl = nf.flows.neural_spline.coupling
    mk = lambda cin, cout: nf.nets.ConvResidualNet(in_channels=cin, out_channels=cout, hidden_channels=16)
    kw = dict(apply_unconditional_transform=True, img_shape=[4, 4]) if uncond else {}
    return cpl.PiecewiseRationalQuadraticCoupling(mask=nf.utils.masks.create_mid_split_binary_mask(2),
                                                  transform_net_create_fn=mk, tails="linear", tail_bound=3.0, **kw)


# Round-2 cases: golden vectors in layers_r2.pt, generated from the unmodified reference by
# `make_golden.py --r2`.  The oracle does not restate these rows; the CUDA product is compared with the reference's
# outputs directly.
LAYER_CASES_R2 = {
    "prq_coupling_img": dict(      # the reference's own test shape [10, 2, 4, 4] (coupling_test.py:49-50)
        build=lambda nf: _img_coupling(nf, False),
        x=lambda: 1.5 * torch.randn(10, 2, 4, 4, generator=_g(41)), scale=0.2),
    "prq_coupling_img_uncond": dict(
        build=lambda nf: _img_coupling(nf, True),
        x=lambda: 1.5 * torch.randn(10, 2, 4, 4, generator=_g(42)), scale=0.2),
    "affine_block_checkerboard": dict(
        build=lambda nf: nf.flows.AffineCouplingBlock(nf.nets.ConvNet2d((3, 8, 6), (3, 3), init_zeros=False),
                                                      split_mode="checkerboard"),
        x=lambda: torch.randn(5, 3, 4, 6, generator=_g(43)), scale=0.1),
    "affine_block_checkerboard_inv": dict(
        build=lambda nf: nf.flows.AffineCouplingBlock(nf.nets.ConvNet2d((3, 8, 6), (3, 3), init_zeros=False),
                                                      split_mode="checkerboard_inv"),
        x=lambda: torch.randn(5, 3, 4, 6, generator=_g(44)), scale=0.1),
    "logit_transform": dict(
        build=lambda nf: nf.transforms.Logit(alpha=0.05),
        x=lambda: torch.rand(6, 3, 8, 8, generator=_g(45)), randomize=False),
}


def _bounded_coupling(nf, uncond):
    cpl = nf.flows.neural_spline.coupling
    mk = lambda cin, cout: nf.nets.ResidualNet(in_features=cin, out_features=cout, hidden_features=16, num_blocks=1)
    return cpl.PiecewiseRationalQuadraticCoupling(mask=nf.utils.masks.create_alternating_binary_mask(6, even=True),
                                                  transform_net_create_fn=mk, num_bins=8,
                                                  apply_unconditional_transform=uncond)       # tails=None (the default)


# tails=None: the bounded spline on [0, 1] (utils/splines.py:88-193 called directly, neural_spline/coupling.py:321-323,
# autoregressive.py:106-108); inputs inside the box.  layers_r3.pt, `make_golden.py --r3`.
LAYER_CASES_R3 = {
    "prq_coupling_bounded": dict(
        build=lambda nf: _bounded_coupling(nf, False),
        x=lambda: torch.rand(64, 6, generator=_g(51)), scale=0.3),
    "prq_coupling_bounded_uncond": dict(
        build=lambda nf: _bounded_coupling(nf, True),
        x=lambda: torch.rand(64, 6, generator=_g(52)), scale=0.3),
    "prq_cdf_bounded": dict(
        build=lambda nf: nf.flows.neural_spline.coupling.PiecewiseRationalQuadraticCDF(shape=[5], num_bins=8),
        x=lambda: torch.rand(40, 5, generator=_g(53)), scale=0.5),
    "marqs_bounded": dict(
        build=lambda nf: nf.flows.neural_spline.autoregressive.MaskedPiecewiseRationalQuadraticAutoregressive(
            features=6, hidden_features=16, num_bins=8, num_blocks=1, init_identity=False),
        x=lambda: torch.rand(48, 6, generator=_g(54)), scale=0.3),
}


def build_cfg2_small(nf, layers=3, d=29, hidden=32):
    flows = []
    for _ in range(layers):
        flows += [nf.flows.CoupledRationalQuadraticSpline(d, 2, hidden), nf.flows.LULinearPermute(d)]
    return nf.NormalizingFlow(nf.distributions.DiagGaussian(d, trainable=False), flows,
                              categoricals=[], catlevels=[], catvdeqs=[])


def build_cfg1(nf, layers=4):
    flows = []
    for _ in range(layers):
        flows += [nf.flows.AffineCouplingBlock(nf.nets.MLP([1, 32, 32, 2], init_zeros=True)),
                  nf.flows.Permute(2, mode="swap")]
    return nf.NormalizingFlow(nf.distributions.DiagGaussian(2), flows, p=nf.distributions.TwoMoons())


def build_cfg3_small(nf, layers=2, d=8, hidden=32):
    flows = []
    for _ in range(layers):
        flows += [nf.flows.AutoregressiveRationalQuadraticSpline(d, 2, hidden), nf.flows.LULinearPermute(d)]
    return nf.NormalizingFlow(nf.distributions.DiagGaussian(d, trainable=False), flows,
                              categoricals=[], catlevels=[], catvdeqs=[])


def build_cfg4_small(nf, d=6, cats=(1, 4), levels=(5, 3), n_main=3, n_vdeq=2):
    """runadultvdeq1ft.py:85-108 recipe, shrunk: one 1-D vdeq per categorical column (shared layers)."""
    vlayers = []
    for i in range(n_vdeq):
        b = torch.tensor([1.0 if i % 2 == 0 else 0.0])
        vlayers += [nf.flows.ActNorm(1), nf.flows.MaskedAffineFlow(b, nf.nets.MLP([1, 4, 1], init_zeros=True),
                                                                    nf.nets.MLP([1, 4, 1], init_zeros=True)),
                    nf.flows.ActNorm(1)]
    vdeqs = [nf.nets.VariationalDequantization(var_flows=vlayers, quants=q) for q in levels]
    flows = []
    for i in range(n_main):
        s, t = _mlp_pair(nf, d, 2 * d)
        flows += [nf.flows.MaskedAffineFlow(_alt_mask(d, i % 2 == 1), t, s), nf.flows.ActNorm(d)]
    q0 = nf.distributions.GaussianMixture(n_modes=4, dim=d)
    return nf.NormalizingFlow(q0, flows, categoricals=list(cats), catlevels=list(levels), catvdeqs=vdeqs)


def cfg4_data(n=64, d=6, cats=(1, 4), levels=(5, 3), seed=31):
    g = _g(seed)
    x = torch.rand(n, d, generator=g)
    for c, q in zip(cats, levels):
        x[:, c] = torch.randint(0, q, (n,), generator=g).float()
    return x


def build_glow_small(nf, L=2, K=2, hidden=16, in_shape=(3, 8, 8), classes=4):
    """example/glow.ipynb:37-73 recipe, shrunk."""
    c, h, w = in_shape
    q0, merges, flows = [], [], []
    for i in range(L):
        flows_ = [nf.flows.GlowBlock(c * 2 ** (L + 1 - i), hidden, split_mode="channel", scale=True)
                  for _ in range(K)]
        flows_ += [nf.flows.Squeeze()]
        flows += [flows_]
        if i > 0:
            merges += [nf.flows.Merge()]
            shape = (c * 2 ** (L - i), h // 2 ** (L - i), w // 2 ** (L - i))
        else:
            shape = (c * 2 ** (L + 1), h // 2 ** L, w // 2 ** L)
        q0 += [nf.distributions.ClassCondDiagGaussian(shape, classes)]
    return nf.MultiscaleFlow(q0, flows, merges)


MODEL_CASES = {
    "cfg2_small": dict(build=build_cfg2_small, x=lambda: 1.5 * torch.randn(128, 29, generator=_g(21)),
                       calls=("forward_kld", "log_prob"), scale=0.1),
    "cfg1": dict(build=build_cfg1, x=lambda: torch.randn(128, 2, generator=_g(22)), calls=("log_prob",),
                 scale=0.1),
    "cfg3_small": dict(build=build_cfg3_small, x=lambda: 1.5 * torch.randn(64, 8, generator=_g(23)),
                       calls=("forward_kld", "log_prob"), scale=0.1),
    "cfg4_small": dict(build=build_cfg4_small, x=cfg4_data, calls=("forward_kld",), dtype=torch.float64,
                       scale=0.2),
    "glow_small": dict(build=build_glow_small,
                       x=lambda: torch.rand(8, 3, 8, 8, generator=_g(24)),
                       y=lambda: torch.randint(0, 4, (8,), generator=_g(25)), calls=("forward_kld",),
                       scale=0.03),
}
