"""GPU parity tests for the tcgen05 kernels (tc_gemm.cu): dense weighted covariance and the dense
Gaussian transform, against the oracle (BLAS restatement of ppi/policies.py:78-93 and
ppi/samplers.py:38-39).  Tolerance 1e-5 norm-wise (BASELINE.json north_star)."""
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


@pytest.fixture(scope="module", params=["tf32x3", "fp16x2"])
def ops(request):
    """Every test runs under both split-product schemes of the tcgen05 kernels (ppi_set_option)."""
    import torch

    assert torch.cuda.is_available()
    from ppi_b200 import _lib, ops as _ops

    scheme = _lib.GEMM_TF32X3 if request.param == "tf32x3" else _lib.GEMM_FP16X2
    _lib.call("ppi_set_option", _lib.OPT_GEMM_SCHEME, scheme)
    assert _lib.query("ppi_get_option", _lib.OPT_GEMM_SCHEME) == scheme
    yield _ops
    _lib.call("ppi_set_option", _lib.OPT_GEMM_SCHEME, _lib.GEMM_FP16X2)


def _problem(K, D, seed, ess_frac=None):
    rng = np.random.default_rng(seed)
    scale = 0.2 + rng.random(D) * 3.0
    x = f32r(1.0 + rng.standard_normal((K, D)) * scale[None, :])
    lw = f32r(-8.0 * rng.random(K))
    nw = f32r(np.exp(lw - logsumexp(lw)))
    return x, lw, nw


@pytest.mark.parametrize("K,D", [(512, 128), (4096, 256), (1000, 200), (3000, 384), (70000, 256),
                                 (8192, 1024), (600, 130)])
def test_wmoments_dense(ops, record, K, D):
    x, lw, nw = _problem(K, D, K + D)
    mu, sigma, ess = O.m_projection(lw, x, naive=False)
    # reference weights in fp64 differ from the fp32 nw by rounding only; use the fp32 ones on both sides
    mu32 = f32r(nw @ x)
    ref = (nw[:, None] * (x - mu32)).T @ (x - mu32)
    cov = host(ops.wmoments_dense(x, nw, shift=mu32))
    assert nerr(cov, ref) < TOL
    assert np.array_equal(cov, cov.T)                        # symmetric by construction
    record(f"wmoments_dense[{K}-{D}]", vs_fp32_weights=nerr(cov, ref), vs_oracle=nerr(cov, sigma))
    assert nerr(cov, sigma) < TOL                            # and equal to the oracle's m_projection
    # linearity: two half batches accumulate to the full one
    h = K // 2
    c2 = ops.wmoments_dense(x[:h], nw[:h], shift=mu32)
    c2 = host(ops.wmoments_dense(x[h:], nw[h:], shift=mu32, cov=c2, accumulate=True))
    assert nerr(c2, ref) < TOL


def test_wmoments_dense_zero_weight_rows(ops):
    K, D = 2048, 256
    x, lw, nw = _problem(K, D, 5)
    nw[::5] = 0.0
    xb = x.copy()
    xb[::5] = np.nan  # failed rollouts: NaN rows with zero weight must not leak
    ref = (nw[:, None] * np.nan_to_num(xb)).T @ np.nan_to_num(xb)
    cov = host(ops.wmoments_dense(xb, nw))
    assert np.isfinite(cov).all() and nerr(cov, ref) < TOL


@pytest.mark.parametrize("K,D", [(256, 256), (1000, 200), (5000, 384), (40000, 256), (2048, 1024),
                                 (300, 129), (70001, 640), (513, 2048)])
def test_sample_gaussian_dense(ops, K, D):
    rng = np.random.default_rng(K * D)
    z = f32r(rng.standard_normal((K, D)))
    a = rng.standard_normal((D, D + 8))
    sigma = a @ a.T / D + 0.2 * np.eye(D)
    L = f32r(np.linalg.cholesky(sigma))
    mu = f32r(rng.standard_normal(D))
    ref = mu[None, :] + z @ L.T
    x = host(ops.sample_gaussian(K, mu, L, z=z))
    assert nerr(x, ref) < TOL
    # Philox mode equals injected mode fed with ppi_philox_normal
    zp = ops.philox_normal(K, D, seed=3, draw_id=1, row_offset=77, order="col")   # the dense sampler's column-quad order
    xa = host(ops.sample_gaussian(K, mu, L, z=zp))
    xb = host(ops.sample_gaussian(K, mu, L, seed=3, draw_id=1, row_offset=77))
    assert np.array_equal(xa, xb)


def test_gaussian_policy_dense_update(ops, record):
    """Config C2 in small (D = 256, K = 8192): sample -> Rastrigin -> ESSPS -> moment match,
    against the oracle on the same injected normals."""
    from ppi_b200 import algorithms as alg, functions as fn, policies as pol, samplers as smp

    K, D = 8192, 256
    rng = np.random.default_rng(0)
    z = np.asarray(rng.standard_normal((K, D)), np.float32)
    smp.InjectedNormals.queue = [z.copy()]
    policy = pol.GaussianPolicy(np.ones(D), 0.5 * np.eye(D), smp.InjectedNormals(D))
    solver = alg.Essps(n_elites=K // 64)
    solver.reset(policy)
    samples, _ = policy(K)
    x32 = np.asarray(samples)
    assert nerr(x32, O.sample_gaussian(z.astype(np.float64), np.ones(D), 0.5 * np.eye(D))) < TOL
    costs = fn.Rastrigin(D)(samples)
    c32 = np.asarray(costs)
    assert nerr(c32, O.rastrigin(x32)) < 2e-6
    solver.update(costs, samples, policy)
    log_w, alpha = O.essps_log_w(c32, K // 64)
    mu, sigma, ess, kl = O.gaussian_weighted_update(np.ones(D), 0.5 * np.eye(D), log_w, x32)
    # (the oracle's m_projection uses the naive einsum: fine at this size)
    tr = solver.trace
    assert abs(tr["alpha"][-1] - alpha) < TOL * alpha
    assert abs(tr["ess"][-1] - ess) < TOL * ess
    assert nerr(policy.mu, mu) < TOL
    assert nerr(policy.sigma, sigma) < TOL
    assert abs(tr["kl"][-1] - kl) < 1e-4 * max(1.0, abs(kl))
    assert abs(tr["ent"][-1] - O.gaussian_entropy(sigma, D)) < 1e-4 * abs(O.gaussian_entropy(sigma, D))
    # second iteration samples from the fitted full covariance
    z2 = np.asarray(rng.standard_normal((K, D)), np.float32)
    smp.InjectedNormals.queue = [z2.copy()]
    x2 = np.asarray(policy(K)[0])
    err2 = nerr(x2, O.sample_gaussian(z2.astype(np.float64), policy.mu, policy.sigma))
    record("gaussian_policy_dense_update second draw", x=err2, cond_sigma=float(np.linalg.cond(policy.sigma)))
    assert err2 < TOL


def test_wmoments_dense_dev_bound_and_overlap(ops):
    """The operand-scale bound may come from the mean pass (ppi_wmoments_diag absdev_max) and may be
    loose; staging on the side stream (several passes, K > 65536) changes nothing bit-wise."""
    import torch

    from ppi_b200 import _lib

    K, D = 150000, 256
    x, lw, nw = _problem(K, D, 3)
    mu_old = f32r(np.ones(D))
    xd, nwd = torch.as_tensor(x, dtype=torch.float32).cuda(), torch.as_tensor(nw, dtype=torch.float32).cuda()
    absdev = torch.zeros(1, dtype=torch.float32, device="cuda")
    s1 = ops.wmoments_diag(xd, nwd, shift1=mu_old, absdev=absdev)
    assert abs(float(absdev.item()) - np.max(np.abs(x - mu_old))) < 1e-6 * np.max(np.abs(x))
    mu32 = f32r(mu_old + host(s1))
    ref = (nw[:, None] * (x - mu32)).T @ (x - mu32)
    base = host(ops.wmoments_dense(xd, nwd, shift=mu32))
    assert nerr(base, ref) < TOL
    for factor in (2.0, 2000.0):
        cov = host(ops.wmoments_dense(xd, nwd, shift=mu32, dev_bound=factor * absdev))
        assert nerr(cov, ref) < TOL
    _lib.call("ppi_set_option", _lib.OPT_GEMM_OVERLAP, 1)
    try:
        overlapped = host(ops.wmoments_dense(xd, nwd, shift=mu32))
    finally:
        _lib.call("ppi_set_option", _lib.OPT_GEMM_OVERLAP, 0)
    assert np.array_equal(overlapped, base)


def test_sample_gaussian_overlap_bitwise(ops):
    import torch

    from ppi_b200 import _lib

    K, D = 140000, 256
    rng = np.random.default_rng(1)
    a = rng.standard_normal((D, D + 8))
    L = f32r(np.linalg.cholesky(a @ a.T / D + 0.2 * np.eye(D)))
    mu = f32r(rng.standard_normal(D))
    xa = ops.sample_gaussian(K, mu, L, seed=9, draw_id=2)
    _lib.call("ppi_set_option", _lib.OPT_GEMM_OVERLAP, 1)
    try:
        xb = ops.sample_gaussian(K, mu, L, seed=9, draw_id=2)
    finally:
        _lib.call("ppi_set_option", _lib.OPT_GEMM_OVERLAP, 0)
    assert torch.equal(xa, xb)
    zp = host(ops.philox_normal(K, D, seed=9, draw_id=2, order="col"))
    assert nerr(host(xa), mu[None, :] + zp @ L.T) < TOL


