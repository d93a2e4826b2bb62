"""GPU parity tests of the reference's GUARD branches and of the iCEM sample re-use, against fixtures
produced by the unmodified reference (tests/golden/make_golden_guards.py):

  * GaussianPolicy, fitted covariance not positive definite -> old sigma + 1e-6 I  (ppi/policies.py:259-268),
    on the SIMT path (D = 8), the x-space tensor-core path and the whitened path (D = 200);
  * kernel policy, Cholesky fails -> mean accepted, covariance_in / sqrt + 1e-6 I, ess = K  (:832-846);
  * feature policy, Cholesky fails -> full revert, ess = K  (:499-515);
  * iCem + Particles on ColouredNoise (ppi/algorithms.py:149-165, ppi/samplers.py:67-102,
    ppi/policies.py:1198-1255), with the particle shift of update_timesteps.

Tolerance 1e-5 norm-wise (north_star); the guard outcomes themselves (which branch ran, ess = K) are exact.
"""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
TOL = 1e-5
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def nerr(a, ref):
    a, ref = np.asarray(a, np.float64), np.asarray(ref, np.float64)
    assert a.shape == ref.shape, (a.shape, ref.shape)
    return np.max(np.abs(a - ref)) / max(np.max(np.abs(ref)), 1e-300)


@pytest.fixture(scope="module")
def g():
    return dict(np.load(os.path.join(ROOT, "tests", "golden", "reference_golden_guards.npz")))


@pytest.fixture(scope="module")
def mods():
    import torch

    assert torch.cuda.is_available()
    from ppi_b200 import algorithms, functions, policies, samplers

    return algorithms, policies, samplers, functions


@pytest.mark.parametrize("tag,route", [("g1", "own"), ("g1", "host"), ("g2", "own"), ("g2", "host")])
def test_gaussian_non_pd_branch(mods, g, tag, route):
    """A rank-deficient sample set (K - 1 < D): np.linalg.cholesky raises in the reference, mu is accepted
    and sigma becomes sigma_old + 1e-6 I.  route "own": the policy's own draw (D = 200: the whitened
    path, whose C = S - m m^T is singular); route "host": the same samples handed back as a host array
    (the x-space path)."""
    alg, pol, smp, fn = mods
    mu0, sig0, z, lw = g[f"{tag}_mu0"], g[f"{tag}_sig0"], g[f"{tag}_z"], g[f"{tag}_lw"]
    K, D = z.shape
    smp.InjectedNormals.queue = [z.copy()]
    policy = pol.GaussianPolicy(mu0.copy(), sig0.copy(), smp.InjectedNormals(D))
    x, _ = policy(K)
    assert nerr(np.asarray(x), g[f"{tag}_x"]) < TOL
    before = policy.n_regularized
    ess, kl = policy.weighted_update(lw, x if route == "own" else g[f"{tag}_x"])
    assert policy.n_regularized == before + 1, "the non-PD branch did not run"
    assert nerr(policy.mu, g[f"{tag}_mu"]) < TOL
    assert nerr(policy.sigma, g[f"{tag}_sigma"]) < TOL
    assert nerr(policy.sigma - sig0, 1e-6 * np.eye(D)) < 1e-3      # exactly the reference's regulariser
    assert abs(ess - float(g[f"{tag}_ess"])) < TOL * float(g[f"{tag}_ess"])
    assert abs(kl - float(g[f"{tag}_kl"])) < 1e-4 * max(1.0, abs(float(g[f"{tag}_kl"])))
    # the policy is still usable: the next draw comes from (mu_new, sigma_old + 1e-6 I)
    smp.InjectedNormals.queue = [z.copy()]
    x2, _ = policy(K)
    ref = g[f"{tag}_mu"][None, :] + z @ np.linalg.cholesky(g[f"{tag}_sigma"]).T
    assert nerr(np.asarray(x2), ref) < TOL


def test_kernel_revert_branch(mods, g):
    alg, pol, smp, fn = mods
    t, xs, lw = g["g4_t"], g["g4_xs"], g["g4_lw"]
    H, d = xs.shape[1:]
    policy = pol.SquaredExponentialKernel(time_sequence=t, action_dimension=d, mean=g["g4_mean_fn"],
                                          covariance_in=np.array([0.7]), covariance_out=g["g4_cout"],
                                          lengthscale=0.08, sampler=smp.MonteCarlo)
    assert nerr(policy.covariance_in, g["g4_cin0"]) < 1e-12
    ess, kl = policy.weighted_update(lw, xs)
    assert ess == xs.shape[0] == int(g["g4_ess"]) and kl == 0.0     # the revert reports ess = K
    assert nerr(policy.mean, g["g4_mean"]) < TOL                      # the mean IS accepted (:817)
    assert nerr(policy.covariance_in, g["g4_cov"]) < 1e-9
    assert nerr(policy.covariance_in_sqrt, g["g4_sqrt"]) < 1e-9      # sqrt + 1e-6 I, not a re-factorisation
    assert nerr(policy.map_sequence, g["g4_map"]) < TOL


def test_features_full_revert_branch(mods, g):
    alg, pol, smp, fn = mods
    ws, lw = g["g5_ws"], g["g5_lw"]
    F, d = ws.shape[1:]
    policy = pol.RbfFeatures(time_sequence=g["g5_t"], action_dimension=d, mean=g["g4_mean_fn"],
                             covariance_in=np.array([0.5]), covariance_out=g["g4_cout"],
                             lengthscale=np.sqrt(3e-2), n_features=F - 1, add_bias=True, use_derivatives=False,
                             sampler=smp.MonteCarlo)
    ess, kl = policy.weighted_update(lw, ws)
    assert ess == ws.shape[0] == int(g["g5_ess"]) and kl == 0.0
    assert np.array_equal(policy.mean, g["g5_mean"])                 # nothing moved: full revert
    assert nerr(policy.covariance_in, g["g5_cov"]) < 1e-12
    # the regular branch right after it
    ess2, _ = policy.weighted_update(g["g5_lw2"], g["g5_ws2"])
    assert abs(ess2 - float(g["g5_ess2"])) < TOL * float(g["g5_ess2"])
    assert nerr(policy.mean, g["g5_mean2"]) < TOL
    assert nerr(policy.covariance_in, g["g5_cov2"]) < TOL


def _tracking_target(t, d):
    amps = np.linspace(-1, 1, d)
    return np.concatenate([amps[k] * np.sin(4 * (k + 1) * t)[:, None] / 2 for k in range(d)], axis=1)


def test_icem_particles_on_coloured_noise(mods, g):
    """iCem stores the best int(0.33 n_elites) SAMPLES in the policy's Particles sampler; the next draw
    overwrites its first rows of noise with them (ppi/samplers.py:92-97), and update_timesteps shifts them
    with the window.  The power-law generator is taken out of the picture on both sides (injected draws;
    mixing matrix = identity): this pins the plumbing around it."""
    alg, pol, smp, fn = mods
    draws, t0, dt = g["g6_draws"], g["g6_t0"], float(g["g6_dt"])
    iters, K, d, H = draws.shape
    policy = pol.ColouredNoise(time_sequence=t0, action_dimension=d, mean=g["g6_mean_fn"],
                               covariance_in=np.array([1.0]), covariance_out=np.diag([0.3, 0.6]),
                               sampler=smp.Particles, beta=2.0,
                               limiter=pol.Limiter(lower=g["g6_lower"], upper=g["g6_upper"]))
    assert isinstance(policy.sampler, smp.Particles) and not policy.sampler.has_particles
    solver = alg.iCem(n_elites=int(g["g6_n_elites"]))
    solver.reset(policy)
    for it in range(iters):
        if it == 2:
            policy.update_timesteps(dt * np.arange(1, H + 1), 1.0)
        policy._mix_cache = ((policy.dim_features, policy.beta), np.eye(policy.dim_features))
        policy.injected_normals = [draws[it].transpose(0, 2, 1).copy()]     # (K, H, d), :1241-1243
        xs, params = policy(K)
        assert nerr(np.asarray(xs), g[f"g6_xs_{it}"]) < TOL, it
        target = _tracking_target(policy.t, d)
        costs = np.mean((np.asarray(xs) - target[None]) ** 2, axis=(1, 2))
        assert nerr(costs, g[f"g6_costs_{it}"]) < TOL
        solver.update(g[f"g6_costs_{it}"], params, policy)               # identical costs into the sort
        assert nerr(policy.mean, g[f"g6_mean_{it}"]) < TOL, it
        assert nerr(policy.std, g[f"g6_std_{it}"]) < TOL, it
        assert policy.sampler.has_particles
        assert nerr(policy.sampler.particles, g[f"g6_particles_{it}"]) < TOL, it
        assert nerr(policy.map_sequence, g[f"g6_map_{it}"]) < TOL, it
    assert nerr(np.asarray(solver.trace["ess"], np.float64), g["g6_ess"]) < TOL
    assert nerr(np.asarray(solver.trace["weight_ent"], np.float64), g["g6_weight_ent"]) < TOL


def test_particles_sampler_protocol(mods):
    """ppi/samplers.py:67-102 as a GaussianPolicy sampler: without particles it is a Monte Carlo draw; once
    particles exist, add_particles unpacks a 3-D shape from the 2-D sample matrix and raises ValueError --
    the reference's behaviour, kept."""
    alg, pol, smp, fn = mods
    D, K = 6, 256
    np.random.seed(5)
    sampler = smp.Particles(D)
    assert not sampler.has_particles and sampler.covariance_scale(10) == 9
    x = sampler(np.zeros(D), np.eye(D), K)
    assert x.shape == (K, D) and abs(x.mean()) < 0.2 and abs(x.std() - 1.0) < 0.1
    z = sampler(np.zeros(D), np.eye(D), K, standard_gaussian=True)
    assert z.shape == (K, D)
    sampler.particles = np.zeros((4, 3, 2))
    assert sampler.has_particles
    with pytest.raises(ValueError):
        sampler(np.zeros(D), np.eye(D), K)
    # iCem on a GaussianPolicy whose sampler is Particles: the elites are stored (algorithms.py:162-164)
    policy = pol.GaussianPolicy(np.ones(D), 0.5 * np.eye(D), smp.Particles(D))
    solver = alg.iCem(n_elites=12)
    solver.reset(policy)
    samples, params = policy(K)
    costs = fn.Rosenbrock(D)(samples)
    solver.update(costs, params, policy)
    order = np.argsort(np.asarray(costs), kind="stable")
    assert policy.sampler.has_particles
    assert np.array_equal(policy.sampler.particles, np.asarray(samples)[order[:solver.n_reuse]])
