"""GPU parity tests of SURVEY 8(f) row 2 (samplers): device-side scrambled-Sobol + erfinv base normals
against scipy.stats.qmc.Sobol -- the engine the reference's QuasiMonteCarlo wraps (ppi/samplers.py:25-44)
-- and the cubature sigma points (:47-64), alone and driving a GaussianPolicy update."""
import numpy as np
import pytest

from oracle import ppi_oracle as O

pytestmark = pytest.mark.gpu
TOL = 1e-5


@pytest.fixture(scope="module")
def mods():
    import torch

    assert torch.cuda.is_available()
    from ppi_b200 import algorithms, functions, policies, samplers

    return algorithms, functions, policies, samplers


def reference_normals(d, seed, calls, shrinkage=0.9999):
    """What ppi/samplers.py:31-37 returns for consecutive calls (standard_gaussian=True)."""
    from scipy.special import erfinv
    from scipy.stats import qmc

    eng = qmc.Sobol(d=d, scramble=True, seed=seed)
    out = []
    for n in calls:
        n_b2 = 2 ** int(np.ceil(np.log(n) / np.log(2)))
        out.append(np.sqrt(2) * erfinv(2 * shrinkage * eng.random(n_b2)[:n] - 1))
    return out


@pytest.mark.parametrize("d,calls", [(3, (100, 300, 7)), (130, (1000, 64)), (2048, (65, 129)), (1, (5000,))])
def test_sobol_normals_match_scipy(mods, d, calls):
    smp = mods[3]
    sampler = smp.QuasiMonteCarlo(d, seed=5)
    for n, ref in zip(calls, reference_normals(d, 5, calls)):
        z, _, _ = sampler.device_normals(n, d)
        z = z.cpu().numpy().astype(np.float64)
        assert z.shape == ref.shape
        assert np.max(np.abs(z - ref)) < 4e-7 * max(1.0, np.max(np.abs(ref))), (d, n)


def test_sobol_rows_are_index_addressable(mods):
    """K-sharding: a rank generating rows [a, b) of a call gets the rows of the full matrix."""
    from ppi_b200 import ops

    smp = mods[3]
    s = smp.QuasiMonteCarlo(17, seed=2)
    sv, sh = s._device_state()
    full = ops.sobol_normal(1000, 17, sv, sh, s.bits, 4096)
    part = ops.sobol_normal(300, 17, sv, sh, s.bits, 4096 + 650)
    assert (full[650:950] == part).all()


def test_sobol_host_protocol(mods):
    smp = mods[3]
    d = 6
    mu, a = np.arange(d, dtype=np.float64), np.random.default_rng(0).standard_normal((d, d + 2))
    sigma = a @ a.T / d + 0.1 * np.eye(d)
    ref = reference_normals(d, 9, (200,))[0]
    x = smp.QuasiMonteCarlo(d, seed=9)(mu, sigma, 200)
    want = mu[None, :] + ref @ np.linalg.cholesky(sigma).T
    assert np.max(np.abs(x - want)) < 1e-5 * np.max(np.abs(want))
    assert smp.QuasiMonteCarlo.covariance_scale(10) == 9


def test_sobol_c_abi_errors(mods):
    import torch
    from ppi_b200 import _lib

    t = torch.zeros(64, dtype=torch.int32, device="cuda")
    o = torch.zeros(64, dtype=torch.float32, device="cuda")
    with pytest.raises(_lib.PpiError):
        _lib.call("ppi_sobol_normal", o.data_ptr(), 4, 2, t.data_ptr(), t.data_ptr(), 33, 0, 0.9999, None)
    with pytest.raises(_lib.PpiError):   # beyond the 2^bits points of the sequence
        _lib.call("ppi_sobol_normal", o.data_ptr(), 4, 2, t.data_ptr(), t.data_ptr(), 30, 2 ** 30 - 2, 0.9999, None)


def test_gaussian_policy_with_qmc(mods):
    alg, fn, pol, smp = mods
    d, K = 20, 512
    rng = np.random.default_rng(1)
    a = rng.standard_normal((d, d + 3))
    sigma = a @ a.T / d + 0.2 * np.eye(d)
    mu = rng.standard_normal(d)
    policy = pol.GaussianPolicy(mu.copy(), sigma.copy(), smp.QuasiMonteCarlo(d, seed=4))
    xs, params = policy(K)
    z = reference_normals(d, 4, (K,))[0]
    want = O.sample_gaussian(z, mu, sigma)
    got = np.asarray(xs)
    assert np.max(np.abs(got - want)) < TOL * np.max(np.abs(want))
    lw = -0.5 * rng.random(K)
    ess, kl = policy.weighted_update(lw, params)
    mu_r, sig_r, ess_r, kl_r = O.gaussian_weighted_update(mu, sigma, lw.astype(np.float32).astype(np.float64),
                                                        want.astype(np.float32).astype(np.float64))
    assert np.max(np.abs(policy.mu - mu_r)) < TOL * np.max(np.abs(mu_r))
    assert np.max(np.abs(policy.sigma - sig_r)) < 5 * TOL * np.max(np.abs(sig_r))
    assert abs(ess - ess_r) < 1e-4 * ess_r


def test_cubature_points(mods):
    alg, fn, pol, smp = mods
    d = 9
    s = smp.CubatureQuadrature(d)
    pts, _, _ = s.device_normals(0, d)
    assert np.array_equal(pts.cpu().numpy(), s.base_points.astype(np.float32))
    assert s.n_samples == 2 * d
    mu, sigma = np.ones(d), 0.5 * np.eye(d) + 0.1
    policy = pol.GaussianPolicy(mu.copy(), sigma.copy(), s)
    xs, params = policy(s.n_samples)
    want = mu[None, :] + s.base_points @ np.linalg.cholesky(sigma).T
    assert np.max(np.abs(np.asarray(xs) - want)) < TOL * np.max(np.abs(want))
    # the sigma points reproduce mean and covariance exactly under uniform weights
    policy.weighted_update(np.zeros(2 * d), params)
    assert np.max(np.abs(policy.mu - mu)) < TOL and np.max(np.abs(policy.sigma - sigma)) < TOL
