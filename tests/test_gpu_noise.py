"""GPU parity tests of the noise priors of SURVEY 8(f) row 2: SmoothExplorationNoise,
SmoothActionNoise (ppi/policies.py:1257-1326) against the reference's own outputs on injected
normals (tests/golden/reference_golden_noise.npz), ColouredNoise (:1198-1254) against the covariance
of the published colorednoise algorithm (the package is not installed: parity unpinned draw by draw),
LinearGaussianDynamicalSystemKernel (:1329-1410) against the reference's Gram builder."""
import os

import numpy as np
import pytest

from oracle import ppi_oracle as O

pytestmark = pytest.mark.gpu
TOL = 1e-5


def nerr(a, ref):
    a, ref = np.asarray(a, np.float64), np.asarray(ref, np.float64)
    assert a.shape == ref.shape, (a.shape, ref.shape)
    return np.max(np.abs(a - ref)) / max(np.max(np.abs(ref)), 1e-300)


@pytest.fixture(scope="module")
def g():
    return np.load(os.path.join(os.path.dirname(__file__), "golden", "reference_golden_noise.npz"))


@pytest.fixture(scope="module")
def pol():
    import torch

    assert torch.cuda.is_available()
    from ppi_b200 import policies

    return policies


@pytest.mark.parametrize("name", ["smooth_exploration", "smooth_action"])
@pytest.mark.parametrize("beta", [0.3, 0.8])
def test_smooth_noise_golden(pol, g, name, beta):
    cls = pol.SmoothExplorationNoise if name == "smooth_exploration" else pol.SmoothActionNoise
    K, H, d = g["z"].shape
    policy = cls(g["t"], d, g["mean"], g["cin"], g["cout"], sampler=None, beta=beta,
                 limiter=pol.Limiter(lower=-1.5 * np.ones(d), upper=1.5 * np.ones(d)))
    policy.injected_normals.append(g["z"])
    xs, params = policy(K)
    tag = f"{name}_{beta}"
    assert nerr(np.asarray(xs), g[tag + "_x"]) < TOL
    ess, kl = policy.weighted_update(g["log_w"], params)
    assert abs(ess - float(g[tag + "_ess"])) < TOL * ess and kl == 0.0
    assert nerr(policy.mean, g[tag + "_mean"]) < TOL
    assert nerr(policy.std, g[tag + "_std"]) < TOL
    assert nerr(policy.map_sequence, g[tag + "_x"][np.argmax(g["log_w"])]) < TOL


def test_smooth_noise_philox_matches_filter(pol):
    """Device Philox draws: the stationary statistics are those of the EMA filter."""
    H, d, K, beta = 16, 2, 200000, 0.4
    policy = pol.SmoothExplorationNoise(0.1 * np.arange(H), d, np.zeros(d), np.array([1.0]), np.eye(d),
                                        sampler=None, beta=beta)
    x = np.asarray(policy(K)[0])[:, :, 0]
    f = pol.ema_filter_matrix(H, beta)
    emp = x.T @ x / K
    assert np.max(np.abs(emp - f @ f.T)) < 0.02


@pytest.mark.parametrize("beta", [0.0, 1.0, 2.0])
def test_coloured_noise_covariance(pol, beta):
    H, d, K = 24, 3, 300000
    std = np.array([0.5, 1.0, 2.0])
    policy = pol.ColouredNoise(0.1 * np.arange(H), d, np.array([0.2, 0.0, -0.1]), np.array([1.0]),
                               np.diag(std ** 2), sampler=None, beta=beta)
    x = np.asarray(policy(K)[0]) - np.array([0.2, 0.0, -0.1])[None, None, :]
    ref = O.powerlaw_psd_gaussian(beta, (200000, H), np.random.default_rng(1))
    ref_cov = ref.T @ ref / ref.shape[0]
    for j in range(d):
        emp = x[:, :, j].T @ x[:, :, j] / K / std[j] ** 2
        assert np.max(np.abs(emp - ref_cov)) < 0.03 * np.max(np.abs(ref_cov)), (beta, j)
    # channels are independent
    cross = x[:, :, 0].T @ x[:, :, 1] / K
    assert np.max(np.abs(cross)) < 0.02
    # the update path is WhiteNoiseIid's
    lw = -2.0 * np.random.default_rng(2).random(K)
    params = policy(K)[1]
    policy.weighted_update(lw, params)
    mean, stdv, ess, _ = O.iid_weighted_update(np.array([0.2, 0.0, -0.1]), lw, np.asarray(params))
    assert nerr(policy.mean, mean) < TOL * 10 and nerr(policy.std, stdv) < TOL * 10


@pytest.mark.parametrize("order", [1, 2, 3])
def test_lgds_kernel(pol, g, order):
    from ppi_b200 import samplers as smp

    d = g["mean"].shape[0]
    policy = pol.LinearGaussianDynamicalSystemKernel(g["t"], d, g["mean"], g["cin"], g["cout"], order,
                                                     smp.MonteCarlo, False)
    assert nerr(policy.covariance_in, g[f"lgds_{order}_cov"]) < 1e-12
    assert nerr(policy.covariance_in_sqrt, np.linalg.cholesky(g[f"lgds_{order}_cov"])) < 1e-6
    xs, params = policy(4096)
    x = np.asarray(xs) - g["mean"][None, None, :]
    emp = np.einsum("bij,blj->il", x, x) / (4096 * np.trace(g["cout"]))
    assert np.max(np.abs(emp - policy.covariance_in)) < 0.15 * np.max(np.abs(policy.covariance_in))
    lw = -2.0 * np.random.default_rng(3).random(4096)
    ess, kl = policy.weighted_update(lw, params)
    assert ess > 1.0 and kl == 0.0
