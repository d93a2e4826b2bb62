"""GPU parity tests of the whitened moment match (ppi_sample_gaussian_keep + ppi_wmoments_white,
tc_gemm.cu): for samples x = mu + L z drawn by the policy itself the M-projection of
ppi/policies.py:78-93 is evaluated on the stored base normals z.  Checked against the oracle's
m_projection on x (fp64) -- tolerance 1e-5 norm-wise (BASELINE.json north_star)."""
import numpy as np
import pytest
from scipy.special import logsumexp

from oracle import ppi_oracle as O

pytestmark = pytest.mark.gpu
TOL = 1e-5


def nerr(a, ref):
    a, ref = np.asarray(a, np.float64), np.asarray(ref, np.float64)
    assert a.shape == ref.shape, (a.shape, ref.shape)
    return np.max(np.abs(a - ref)) / max(np.max(np.abs(ref)), 1e-300)


def host(t):
    return t.detach().cpu().numpy().astype(np.float64)


def f32r(x):
    return np.asarray(x, np.float32).astype(np.float64)


@pytest.fixture(scope="module")
def ops():
    import torch

    assert torch.cuda.is_available()
    from ppi_b200 import _lib, ops as _ops

    _lib.call("ppi_set_option", _lib.OPT_GEMM_SCHEME, _lib.GEMM_FP16X2)
    return _ops


def _factor(D, rng):
    a = rng.standard_normal((D, D + 8))
    return f32r(np.linalg.cholesky(a @ a.T / D + 0.2 * np.eye(D))), f32r(rng.standard_normal(D))


def _weights(K, rng, spread=8.0):
    lw = f32r(-spread * rng.random(K))
    return lw, f32r(np.exp(lw - logsumexp(lw)))


@pytest.mark.parametrize("K,D", [(512, 256), (1000, 200), (3000, 384), (600, 130), (8192, 1024),
                                 (150000, 256), (70001, 640)])
@pytest.mark.parametrize("injected", [False, True])
def test_wmoments_white_ops(ops, record, K, D, injected):
    rng = np.random.default_rng(K + D)
    L, mu = _factor(D, rng)
    lw, nw = _weights(K, rng)
    if injected:
        z = f32r(rng.standard_normal((K, D)))
        x, base = ops.sample_gaussian_keep(K, mu, L, z=z)
        ref_x = ops.sample_gaussian(K, mu, L, z=z)
    else:
        z = host(ops.philox_normal(K, D, seed=5, draw_id=3, row_offset=11, order="col"))
        x, base = ops.sample_gaussian_keep(K, mu, L, seed=5, draw_id=3, row_offset=11)
        ref_x = ops.sample_gaussian(K, mu, L, seed=5, draw_id=3, row_offset=11)
    import torch

    assert torch.equal(x, ref_x)                               # the store changes nothing in x
    import torch as _t

    s2 = _t.zeros(D, dtype=_t.float64, device="cuda")
    s1, s = ops.wmoments_white(base, nw, s2=s2)
    m_ref = nw @ z
    s_ref = (nw[:, None] * z).T @ z
    assert nerr(host(s1), m_ref) < TOL
    assert nerr(host(s2), np.diag(s_ref)) < 2e-7        # the round-to-nearest diagonal (accurate trace for the KL)
    cov = host(s)
    assert nerr(cov, s_ref) < TOL
    assert np.array_equal(cov, cov.T)
    # whitened identities against the M-projection of x itself (ppi/policies.py:84-92)
    mu_x, sigma_x, _ = O.m_projection(lw, host(x), naive=False)
    m = host(s1)
    e_mu, e_sig = nerr(mu + L @ m, mu_x), nerr(L @ (cov - np.outer(m, m)) @ L.T, sigma_x)
    record(f"wmoments_white_ops[{K}-{D}-{'injected' if injected else 'philox'}]", m=nerr(host(s1), m_ref),
           S=nerr(cov, s_ref), mu_x=e_mu, sigma_x=e_sig)
    assert e_mu < TOL
    assert e_sig < TOL
    # linearity: accumulate on top of an existing partial
    s1b, sb = ops.wmoments_white(base, 0.25 * nw)
    s1b, sb = ops.wmoments_white(base, 0.75 * nw, s1=s1b, cov=sb, accumulate=True)
    assert nerr(host(s1b), m_ref) < TOL and nerr(host(sb), s_ref) < TOL


def test_wmoments_white_zero_and_one_hot(ops):
    """Rows with zero weight (failed rollouts) contribute nothing; a one-hot weight vector gives the
    outer product of that base-normal row (the alpha -> infinity limit)."""
    K, D = 4096, 256
    rng = np.random.default_rng(2)
    L, mu = _factor(D, rng)
    z = host(ops.philox_normal(K, D, seed=1, order="col"))
    _, base = ops.sample_gaussian_keep(K, mu, L, seed=1)
    _, nw = _weights(K, rng, spread=30.0)   # 13 decades of dynamic range
    nw[::3] = 0.0
    nw /= nw.sum()
    nw = f32r(nw)
    s1, s = ops.wmoments_white(base, nw)
    assert nerr(host(s1), nw @ z) < TOL and nerr(host(s), (nw[:, None] * z).T @ z) < TOL
    hot = np.zeros(K)
    hot[1234] = 1.0
    s1, s = ops.wmoments_white(base, hot)
    assert nerr(host(s1), z[1234]) < TOL and nerr(host(s), np.outer(z[1234], z[1234])) < TOL


@pytest.mark.parametrize("philox", [False, True])
def test_gaussian_policy_white_update(ops, record, philox):
    """Two iterations of sample -> Rastrigin -> ESSPS -> moment match through the public classes at
    D = 384; the fitted mu / sigma / KL / entropy equal the oracle's update computed from x."""
    from ppi_b200 import algorithms as alg, functions as fn, policies as pol, samplers as smp

    K, D = 16384, 384
    rng = np.random.default_rng(0)
    if philox:
        sampler = smp.MonteCarlo(D, seed=77)
    else:
        smp.InjectedNormals.queue = [np.asarray(rng.standard_normal((K, D)), np.float32) for _ in range(2)]
        sampler = smp.InjectedNormals(D)
    policy = pol.GaussianPolicy(np.ones(D), 0.5 * np.eye(D), sampler)
    solver = alg.Essps(n_elites=K // 16)
    solver.reset(policy)
    for it in range(2):
        mu_old, sigma_old = policy.mu.copy(), policy.sigma.copy()
        samples, params = policy(K)
        costs = fn.Rastrigin(D)(samples)      # Philox: the draw is launched here, cost fused, x not written
        assert getattr(samples, "white", None) is not None
        x32, c32 = np.asarray(samples), np.asarray(costs)
        solver.update(costs, params, policy)
        log_w, alpha = O.essps_log_w(c32, K // 16)
        mu, sigma, ess, kl = O.gaussian_weighted_update(mu_old, sigma_old, log_w, x32, naive=False)
        assert abs(solver.trace["alpha"][-1] - alpha) < TOL * alpha
        assert abs(solver.trace["ess"][-1] - ess) < TOL * ess
        record(f"gaussian_policy_white_update[{'philox' if philox else 'injected'}] it{it}",
               mu=nerr(policy.mu, mu), sigma=nerr(policy.sigma, sigma))
        assert nerr(policy.mu, mu) < TOL, it
        assert nerr(policy.sigma, sigma) < TOL, it
        assert abs(solver.trace["kl"][-1] - kl) < 1e-4 * max(1.0, abs(kl))
        ent = O.gaussian_entropy(sigma, D)
        assert abs(solver.trace["ent"][-1] - ent) < 1e-4 * abs(ent)


def test_white_handle_goes_stale(ops, record):
    """A handle drawn before the policy changed must not take the whitened path (its z no longer
    maps to x through the current mu / L): the generic path gives the same answer as a host copy."""
    from ppi_b200 import policies as pol, samplers as smp

    K, D = 4096, 256
    rng = np.random.default_rng(4)
    policy = pol.GaussianPolicy(np.zeros(D), np.eye(D), smp.MonteCarlo(D, seed=3))
    samples, _ = policy(K)
    lw = f32r(-4.0 * rng.random(K))
    policy.mu = np.ones(D)                                    # state change after the draw
    policy.weighted_update(lw, samples)
    mu, sigma, _ = O.m_projection(lw, np.asarray(samples), naive=False)
    assert nerr(policy.mu, mu) < TOL and nerr(policy.sigma, sigma) < TOL
    # and a handle reused after its own update is stale as well
    samples2, _ = policy(K)
    policy.weighted_update(lw, samples2)
    first = policy.sigma.copy()
    policy.weighted_update(lw, samples2)
    mu2, sigma2, _ = O.m_projection(lw, np.asarray(samples2), naive=False)
    record("white_handle_goes_stale", second=nerr(policy.sigma, sigma2), first=nerr(first, sigma2))
    assert nerr(policy.sigma, sigma2) < TOL and nerr(first, sigma2) < TOL


@pytest.mark.parametrize("trigger", ["sample", "update"])
def test_prefetched_normals_are_the_same_draws(ops, trigger):
    """Base normals generated ahead of the sampling call on a side stream (ppi_philox_prefetch +
    ppi_sample_gaussian_pre): the same Philox matrix, hence bit-identical samples and updates."""
    from ppi_b200 import policies as pol, samplers as smp

    K, D = 3000, 384          # ragged: K is not a multiple of the 256-row GEMM block
    rng = np.random.default_rng(8)
    lws = [f32r(-3.0 * rng.random(K)) for _ in range(4)]

    def run(prefetch):
        policy = pol.GaussianPolicy(0.3 * np.ones(D), 0.5 * np.eye(D) + 0.01, smp.MonteCarlo(D, seed=11))
        policy.prefetch_normals = prefetch
        policy.prefetch_min_elements = 0
        xs, used = [], []
        for lw in lws:
            used.append(policy._pre is not None)      # a prefetched draw is waiting for this call
            samples, _ = policy(K)
            xs.append(np.asarray(samples).copy())
            policy.weighted_update(lw, samples)
        return xs, policy.mu.copy(), policy.sigma.copy(), used

    xs0, mu0, sig0, _ = run(None)
    xs1, mu1, sig1, used = run(trigger)
    assert any(used)
    for a, b in zip(xs0, xs1):
        assert np.array_equal(a, b)
    assert np.array_equal(mu0, mu1) and np.array_equal(sig0, sig1)


def test_prefetch_recycles_a_store_only_after_invalidating_its_handle(ops):
    """Two draws without an update in between: the older handle's base-normal store is recycled by the
    next prefetch, so its update must take the x-space path -- and still be right."""
    from ppi_b200 import policies as pol, samplers as smp

    K, D = 2048, 256
    rng = np.random.default_rng(9)
    policy = pol.GaussianPolicy(np.zeros(D), np.eye(D), smp.MonteCarlo(D, seed=5))
    policy.prefetch_normals = "sample"
    policy.prefetch_min_elements = 0
    policy.defer_draws = False                          # eager draws: every call launches and prefetches the next
    handles = [policy(K)[0] for _ in range(4)]          # draws 1.. come from prefetched stores
    assert not handles[1].white[2].valid and handles[3].white[2].valid
    lw = f32r(-4.0 * rng.random(K))
    x1 = np.asarray(handles[1]).copy()
    policy.weighted_update(lw, handles[1])
    mu, sigma, _ = O.m_projection(lw, x1, naive=False)
    assert nerr(policy.mu, mu) < TOL and nerr(policy.sigma, sigma) < TOL


def test_prefetch_c_abi_errors(ops):
    import torch
    from ppi_b200 import _lib

    buf = torch.zeros(1 << 20, dtype=torch.uint8, device="cuda")
    p = buf.data_ptr() + (-buf.data_ptr()) % 1024
    with pytest.raises(_lib.PpiError):     # store far too small for K x D
        _lib.call("ppi_philox_prefetch", 1, 0, 0, 65536, 2048, p, 1 << 19, None)
    with pytest.raises(_lib.PpiError):     # misaligned store
        _lib.call("ppi_philox_prefetch", 1, 0, 0, 256, 256, p + 8, 1 << 19, None)
    with pytest.raises(_lib.PpiError):     # D <= 128 has no tensor-core path
        _lib.call("ppi_philox_prefetch", 1, 0, 0, 256, 64, p, 1 << 19, None)


@pytest.mark.parametrize("name", ["Rastrigin", "Styblinski"])
@pytest.mark.parametrize("K,D", [(3000, 384), (70001, 640), (4096, 2048), (600, 130)])
def test_fused_cost_in_the_sampling_epilogue(ops, record, name, K, D):
    """A Philox draw of a GaussianPolicy is deferred until its first use: `f(samples)` with a separable test
    function launches the sampler with the cost in its epilogue and never writes x (ppi/algorithms.py:85-86
    fused).  The costs equal the stand-alone cost kernel's on the re-drawn x, x equals the eager draw bit for
    bit, and the whitened update that follows is bit-identical to the one of an eager draw."""
    import torch
    from ppi_b200 import functions as fn, policies as pol, samplers as smp
    from ppi_b200.device import DeferredGaussianSamples

    rng = np.random.default_rng(K + D)
    a = rng.standard_normal((D, D + 8))
    sigma = a @ a.T / D + 0.2 * np.eye(D)
    mu = 0.5 * rng.standard_normal(D)
    f = getattr(fn, name)(D)
    lw_seed = f32r(-3.0 * rng.random(K))

    def run(defer):
        policy = pol.GaussianPolicy(mu.copy(), sigma.copy(), smp.MonteCarlo(D, seed=31))
        policy.defer_draws = defer
        samples, params = policy(K)
        assert isinstance(samples, DeferredGaussianSamples) == defer
        costs = f(samples)
        launched_x = getattr(samples, "materialised", True)
        c = costs.tensor.clone()
        x = samples.tensor.clone()
        policy.weighted_update(lw_seed, samples)
        return c, x, policy.mu.copy(), policy.sigma.copy(), launched_x

    c1, x1, mu1, sig1, mat1 = run(True)
    c0, x0, mu0, sig0, mat0 = run(False)
    assert mat1 is False and mat0 is True           # the fused launch did not write x
    assert torch.equal(x1, x0)                      # the re-drawn samples are the eager ones
    ref = {"Rastrigin": O.rastrigin, "Styblinski": O.styblinski}[name](host(x0))
    e_fused, e_kernel = nerr(host(c1), ref), nerr(host(c0), ref)
    record(f"fused_cost[{name}-{K}-{D}]", fused=e_fused, kernel=e_kernel)
    assert e_fused < 2e-6 and e_kernel < 2e-6
    assert np.array_equal(mu1, mu0) and np.array_equal(sig1, sig0)


def test_deferred_draws_with_prefetched_stores(ops):
    """Deferred draws + prefetch: the store of draw t+1 is generated on a side stream behind update t; the deferred
    handle of draw t+1 consumes it (fused cost, no x), a handle whose store slot is recycled before it ran
    generates its own.  Everything bit-identical to the run without prefetch."""
    from ppi_b200 import functions as fn, policies as pol, samplers as smp

    K, D = 3000, 384
    rng = np.random.default_rng(12)
    lws = [f32r(-3.0 * rng.random(K)) for _ in range(4)]
    f = fn.Rastrigin(D)

    def run(prefetch):
        policy = pol.GaussianPolicy(0.3 * np.ones(D), 0.5 * np.eye(D) + 0.01, smp.MonteCarlo(D, seed=21))
        policy.prefetch_normals = prefetch
        policy.prefetch_min_elements = 0
        costs, used = [], []
        for lw in lws:
            samples, _ = policy(K)
            used.append(getattr(samples, "_pre", None) is not None)
            costs.append(host(f(samples).tensor))
            assert not samples.materialised
            policy.weighted_update(lw, samples)
        # two draws without an update: the first one's prefetched store may be recycled before it runs
        h1, h2 = policy(K)[0], policy(K)[0]
        x1, x2 = host(h1.tensor), host(h2.tensor)
        return costs, policy.mu.copy(), policy.sigma.copy(), used, x1, x2

    c0, mu0, sig0, used0, xa0, xb0 = run(None)
    for mode in ("update", "sample"):
        c1, mu1, sig1, used1, xa1, xb1 = run(mode)
        # without prefetch only the FIRST draw of the fresh policy finds a store: its generator ran beside the
        # upload and factorisation of the host sigma (GaussianPolicy.__call__)
        assert used0[0] and not any(used0[1:]) and any(used1[1:]), (mode, used0, used1)
        for a, b in zip(c0, c1):
            assert np.array_equal(a, b)
        assert np.array_equal(mu0, mu1) and np.array_equal(sig0, sig1)
        assert np.array_equal(xa0, xa1) and np.array_equal(xb0, xb1)
