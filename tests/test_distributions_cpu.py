"""Base distributions / targets of the run scripts beyond SURVEY row f1 (vcnf_b200/distributions/extra.py,
target.NealsFunnel): plain library ops, checked against closed forms on the CPU (the reference hard-codes device='cuda'
in their constructors, so it cannot be instantiated in the GPU-less container)."""
import math

import torch
from torch import distributions as D

import vcnf_b200 as nf


def test_names_are_reachable_like_in_the_run_scripts():
    for n in ("MultivariateGaussian", "T", "GGD", "GMM", "GaussianMixture", "DiagGaussian"):
        assert hasattr(nf.distributions.base, n), n
    assert hasattr(nf.distributions.target, "NealsFunnel")


def test_multivariate_gaussian_covariance_and_shapes():
    q = nf.distributions.base.MultivariateGaussian(n_dim=3, trainable=False, scale=1.0)
    assert torch.equal(q.scale.cpu(), torch.eye(3, dtype=torch.double) + 1.0)      # identity + scale in every entry
    z, lp = q(7)
    assert z.shape == (7, 3) and lp.shape == (7,)
    want = D.MultivariateNormal(torch.zeros(3, dtype=torch.double), torch.eye(3, dtype=torch.double) + 1.0).log_prob(z.cpu())
    assert torch.allclose(q.log_prob(z).cpu(), want)


def test_student_t_is_the_feature_mean_and_df_quirk():
    q = nf.distributions.base.T(n_dim=4, df=3., trainable=False)
    assert torch.equal(q.df.cpu(), torch.full((4,), 4.0, dtype=torch.double))      # non-trainable: ones + df
    qt = nf.distributions.base.T(n_dim=4, df=3., trainable=True)
    assert torch.equal(qt.df.detach().cpu(), torch.full((4,), 3.0, dtype=torch.double))
    z = torch.randn(5, 4, dtype=torch.double, device=q.df.device)
    want = D.StudentT(torch.full((4,), 4.0, dtype=torch.double)).log_prob(z.cpu()).mean(1)
    assert torch.allclose(q.log_prob(z).cpu(), want)
    zs, lp = q(6)
    assert zs.shape == (6, 4) and lp.shape == (6,)


def test_ggd_density_is_gaussian_and_samples_are_symmetric():
    torch.manual_seed(0)
    q = nf.distributions.base.GGD(n_dim=2, beta=2., trainable=False)
    z = torch.randn(9, 2, dtype=torch.double, device=q.loc.device)
    want = D.Normal(0., 1.).log_prob(z.cpu()).mean(1)
    assert torch.allclose(q.log_prob(z).cpu(), want)
    zs, lp = q(4000)
    assert zs.shape == (4000, 2) and lp.shape == (4000,) and torch.isfinite(zs).all()
    assert abs(float(zs.mean())) < 0.1                                           # random sign: symmetric about loc


def test_gmm_grid_and_log_prob():
    dev = "cuda" if torch.cuda.is_available() else "cpu"
    w = torch.ones(4, dtype=torch.double, device=dev) / 4
    q = nf.distributions.base.GMM(weights=w, mbase=torch.tensor(2.0, dtype=torch.double, device=dev),
                                  vbase=torch.tensor(5.0, dtype=torch.double, device=dev),
                                  scale=torch.tensor(1.5, dtype=torch.double, device=dev), n_cell=4, dim=3)
    grid = torch.arange(1, 5, dtype=torch.double)
    assert torch.allclose(q.trsf_gridm().cpu()[:, 0], torch.log(1.5 * grid) / math.log(2.0))
    z, lp = q(11)
    assert z.shape == (11, 3) and lp.shape == (11,)
    assert torch.allclose(q.log_prob(z), lp)


def test_neals_funnel_matches_the_two_normals():
    t = nf.distributions.target.NealsFunnel(v1shift=0.5, v2shift=0.1)
    z = torch.randn(13, 2)
    want = D.Normal(0., 1.5).log_prob(z[:, 0]) + D.Normal(torch.zeros(13), torch.exp(0.5 * z[:, 0]) + 0.1).log_prob(z[:, 1])
    assert torch.allclose(t.log_prob(z), want, atol=1e-6)
