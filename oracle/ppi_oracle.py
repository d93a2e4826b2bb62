"""
CPU oracle for the PPI update path -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Plain NumPy/SciPy fp64 restatement of the reference algorithm (the reference is itself
pure NumPy/SciPy, so the restatement is NumPy too).  Every function cites the reference
file:line (relative to /root/reference) that it follows.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference``
legs may import this module, and only as the checker or the timed CPU baseline.  The product
package (``monte-carlo-posterior-policy-iteration_b200/ppi_b200``) never imports it and has no CPU fallback.

Pinning: the reference ships no tests or golden vectors, so this oracle is
pinned against OUTPUTS OF THE REFERENCE ITSELF: ``tests/golden/make_golden.py`` imports
``/root/reference/ppi`` unmodified (three compatibility shims), runs it on seeded
injected inputs and commits the results under ``tests/golden/*.npz``;
``tests/test_oracle_golden.py`` checks every function here against those fixtures
(and against the live reference when ``/root/reference`` is present).

Third-party arithmetic on the path (not vendored in the reference): NumPy (pinned 1.21.5),
SciPy (pinned 1.7.3: ``optimize.minimize`` L-BFGS-B, ``optimize.minimize_scalar`` Brent,
``special.logsumexp``) -- see ppi.yaml:10-14.  They are called here exactly as the reference calls them.
"""
from __future__ import annotations

import numpy as np
from scipy.optimize import minimize, minimize_scalar
from scipy.special import logsumexp

EPS = np.finfo(np.float64).tiny  # ppi/algorithms.py:32, ppi/policies.py:36
SIGMA_MIN = 1e-6  # ppi/policies.py:37
ALPHA_LOWER, ALPHA_UPPER = 1e-5, 5e2  # ppi/algorithms.py:36-37


# ----------------------------------------------------------------------------------------------
# a2  NaN filter                                                        ppi/algorithms.py:57-68
# ----------------------------------------------------------------------------------------------
def filter_nan(costs, weights):
    bad = np.isnan(costs)
    if bad.all():
        return np.zeros_like(costs), weights
    good = ~bad
    return costs[good], weights[good, ...]


# ----------------------------------------------------------------------------------------------
# a10 cost normalisation                      ppi/algorithms.py:184,510,541,597 ; MPPI :472
# ----------------------------------------------------------------------------------------------
def normalise(costs):
    return (costs - np.min(costs)) / (np.max(costs) - np.min(costs) + EPS)


def ess_of(log_w):
    """exp(-LSE(2 log nw))                                   ppi/policies.py:84-87"""
    log_nw = log_w - logsumexp(log_w)
    return np.exp(-logsumexp(2 * log_nw))


# ----------------------------------------------------------------------------------------------
# a11-a15 temperature / weights
# ----------------------------------------------------------------------------------------------
def reps_dual(alpha, costs_, epsilon):
    """ppi/algorithms.py:186-191 (no max shift: costs_ in [0,1])."""
    w_ = np.exp(-alpha * costs_)
    return epsilon / alpha + np.log(w_.mean()) / alpha


def reps_dual_grad(alpha, costs_, epsilon):
    """ppi/algorithms.py:196-203."""
    w_ = np.exp(-alpha * costs_)
    a2 = alpha ** -2
    ndg = a2 * (epsilon + np.log(w_.mean())) + np.einsum("b,b->", w_, costs_) / (
        alpha * w_.sum()
    )
    return -ndg


def reps_log_w(costs, epsilon):
    """Reps.update up to the weights                         ppi/algorithms.py:183-215."""
    costs_ = normalise(costs)
    res = minimize(
        lambda a: reps_dual(a.item(), costs_, epsilon),
        jac=lambda a: np.array([reps_dual_grad(a.item(), costs_, epsilon)]),
        method="L-BFGS-B",
        bounds=((ALPHA_LOWER, ALPHA_UPPER),),
        x0=np.array([1.0]),
        options={"maxiter": 1000},
    )
    alpha = res.x
    return -alpha * costs_, alpha


def reps_kl(alpha, costs_):
    """KL(nw || uniform) of the Gibbs weights; stationarity of the dual <=> KL = epsilon."""
    log_w = -alpha * costs_
    log_nw = log_w - logsumexp(log_w)
    return float((np.exp(log_nw) * (log_nw + np.log(costs_.shape[0]))).sum())


def essps_log_w(costs, n_elites):
    """Essps.update up to the weights                        ppi/algorithms.py:596-615.
    Brent without bounds = the effective algorithm under the pinned SciPy 1.7.3."""
    costs_ = normalise(costs)
    target = int(n_elites)

    def ess_err(alpha):
        return np.abs(ess_of(-alpha * costs_) - target)

    res = minimize_scalar(ess_err, method="brent", options={"maxiter": 5000})
    return -res.x * costs_, res.x


def lbps_objective(alpha, costs_, delta):
    """ppi/algorithms.py:543-555."""
    lambda_ = np.sqrt((1 - delta) / delta)
    log_w_ = -alpha * costs_
    log_nw_ = log_w_ - logsumexp(log_w_)
    nw_ = np.exp(log_nw_)
    ess = np.exp(-logsumexp(2 * log_nw_))
    return np.einsum("b,b->", nw_, costs_) + lambda_ / np.sqrt(ess)


def lbps_log_w(costs, delta):
    """Lbps.log_w                                            ppi/algorithms.py:539-567."""
    costs_ = normalise(costs)
    res = minimize_scalar(
        lambda a: lbps_objective(a, costs_, delta),
        method="brent",
        options={"maxiter": 5000},
    )
    return -res.x * costs_, res.x


def ais_log_w(costs, alpha):
    """Ais.log_w                                             ppi/algorithms.py:509-511."""
    return -normalise(costs) * alpha, alpha


def mppi_log_w(costs, alpha):
    """MppiBase.update                                       ppi/algorithms.py:472-473."""
    return -(costs - np.min(costs)) * alpha, alpha


def cem_log_w(costs, n_elites):
    """Cem.update                                            ppi/algorithms.py:118-120.
    Ties at the k-th cost are resolved by (cost, index) -- np.argsort's introsort is unstable
    (SURVEY 7.3 item 6); tests use tie-free costs."""
    idxs = np.lexsort((np.arange(costs.shape[0]), costs))[:n_elites]
    log_w = -1e12 * np.ones_like(costs)
    log_w[idxs] = 0.0
    return log_w, idxs


def weight_entropy(log_w):
    """sum nw log nw                                         ppi/algorithms.py:123,127,621-622."""
    log_nw = log_w - logsumexp(log_w)
    return (log_nw * np.exp(log_nw)).sum()


# ----------------------------------------------------------------------------------------------
# a16-a21 moment matching
# ----------------------------------------------------------------------------------------------
def m_projection(log_w, samples, naive=True):
    """ppi/policies.py:78-93.  naive=True keeps the reference's einsum (its CPU cost);
    naive=False is the BLAS restatement of the same math ("optimised CPU")."""
    lse_w = logsumexp(log_w)
    log_nw = log_w - lse_w
    nw = np.exp(log_nw)
    ess = np.exp(-logsumexp(2 * log_nw))
    mu = np.einsum("b, bi->i", nw, samples)
    dist = samples - mu[None, :]
    if naive:
        sigma = np.einsum("b,bi,bj->ij", nw, dist, dist)
    else:
        sigma = (nw[:, None] * dist).T @ dist
    sigma = 0.5 * (sigma + sigma.T)
    return mu, sigma, ess


def m_projection_mavn(log_w, samples, covariance_in, covariance_out, naive=True, iterations=1,
                      update_out=False):
    """ppi/policies.py:96-143: matrix-normal MLE; `iterations` rounds of the flip-flop update, the output
    covariance re-estimated only when update_out (both use the DIAGONAL of the other factor, :123,128)."""
    assert iterations > 0
    d_in, d_out = samples.shape[1], covariance_out.shape[0]
    lse_w = logsumexp(log_w)
    log_nw = log_w - lse_w
    nw = np.exp(log_nw)
    ess = np.exp(-logsumexp(2 * log_nw))
    mu = np.einsum("b, bij->ij", nw, samples)
    diff = samples - mu[None, ...]
    for _ in range(iterations):
        out_inv = np.diag(1.0 / np.diag(covariance_out))
        if naive:
            cov_in = np.einsum("b,bij,jk,blk->il", nw, diff, out_inv, diff) / d_out
        else:
            s = diff * np.sqrt(nw[:, None, None] / np.diag(covariance_out)[None, None, :])
            a = s.transpose(1, 0, 2).reshape(s.shape[1], -1)
            cov_in = a @ a.T / d_out
        cov_in = 0.5 * (cov_in + cov_in.T)
        if update_out:
            in_inv = np.diag(1.0 / np.diag(cov_in))
            covariance_out = np.einsum("b,bij,ik,bkl->jl", nw, diff, in_inv, diff) / d_in   # :131-133
            covariance_out = 0.5 * (covariance_out + covariance_out.T)
    return mu, cov_in, covariance_out, ess


def vec(x):
    """ppi/policies.py:161-171: column-major vectorisation."""
    if x.ndim == 3:
        a, b, c = x.shape
        return x.reshape((a, b * c), order="F")
    return x.reshape((-1, 1), order="F")


def matrix_gaussian_kl(mean_1, cov_in_1, cov_out_1, mean_2, cov_in_2, cov_out_2):
    """ppi/policies.py:174-210 (output factors normalised to trace p before the determinants)."""
    eps = np.finfo(np.float64).tiny
    n, p = mean_1.shape
    diff = mean_2 - mean_1
    sf1, sf2 = p / np.trace(cov_out_1), p / np.trace(cov_out_2)
    cov_out_1, cov_out_2 = cov_out_1 * sf1, cov_out_2 * sf2
    cov_in_1, cov_in_2 = cov_in_1 / sf1, cov_in_2 / sf2
    return (0.5 * (n * np.log(max(eps, np.linalg.det(cov_out_2))) - n * np.log(max(eps, np.linalg.det(cov_out_1)))
                   + p * np.log(max(eps, np.linalg.det(cov_in_2))) - p * np.log(max(eps, np.linalg.det(cov_in_1)))
                   + np.trace(np.kron(np.linalg.solve(cov_out_2, cov_out_1), np.linalg.solve(cov_in_2, cov_in_1)))
                   + vec(diff).T @ vec(np.linalg.solve(cov_in_2, np.linalg.solve(cov_out_2, diff.T).T))
                   - n * p)).item()


def gaussian_kl(mu_1, sigma_1, mu_2, sigma_2):
    """ppi/policies.py:146-158."""
    d = sigma_1.shape[0]
    diff = mu_2 - mu_1
    return 0.5 * (
        np.log(max(EPS, np.linalg.det(sigma_2)))
        - np.log(max(EPS, np.linalg.det(sigma_1)))
        + np.trace(np.linalg.solve(sigma_2, sigma_1))
        + diff @ np.linalg.solve(sigma_2, diff)
        - d
    )


def gaussian_entropy(sigma, d):
    """ppi/policies.py:213-216."""
    return 0.5 * np.log(max(EPS, np.linalg.det(sigma))) + (d / 2) * (1 + np.log(2 * np.pi))


def gaussian_weighted_update(mu, sigma, log_w, samples, diagonal=False, update_cov=True, naive=True):
    """GaussianPolicy.weighted_update                        ppi/policies.py:251-269.
    Returns (mu', sigma', ess, kl).  naive: see m_projection."""
    mu_, sigma_, ess = m_projection(log_w, samples, naive=naive)
    if diagonal:
        sigma_ = np.diag(np.diag(sigma_))
    new_sigma = sigma.copy()
    try:
        if update_cov:
            np.linalg.cholesky(sigma_)
            new_sigma = sigma_
    except np.linalg.LinAlgError:
        new_sigma = sigma + SIGMA_MIN * np.eye(mu.shape[0])
    return mu_, new_sigma, ess, gaussian_kl(mu_, new_sigma, mu, sigma)


def kernel_weighted_update(mean_fn, cov_in, cov_in_sqrt, cov_out, log_w, samples, update_cov=True):
    """BaseKernel.weighted_update                            ppi/policies.py:804-848.
    Returns (mean', cov_in', cov_in_sqrt', ess, map_sequence)."""
    map_sequence = samples[np.argmax(log_w), ...]
    corrected = samples - mean_fn[None, None, :]
    mean_, cov_in_, _, ess = m_projection_mavn(log_w, corrected, cov_in, cov_out)
    try:
        if update_cov:
            sqrt_ = np.linalg.cholesky(cov_in_)
            return mean_, cov_in_, sqrt_, ess, map_sequence
        return mean_, cov_in, cov_in_sqrt, ess, map_sequence
    except np.linalg.LinAlgError:
        h = cov_in.shape[0]
        return (mean_, cov_in + SIGMA_MIN * np.eye(h), cov_in_sqrt + SIGMA_MIN * np.eye(h),
                samples.shape[0], map_sequence)


def features_weighted_update(mean, cov_in, cov_in_sqrt, cov_out, log_w, ws, update_cov=True):
    """BaseFeatures.weighted_update                          ppi/policies.py:461-516.
    Returns (mean', cov_in', cov_in_sqrt', ess)."""
    mean_, cov_in_, _, ess = m_projection_mavn(log_w, ws, cov_in, cov_out)
    try:
        cov_in_ = cov_in_ + 1e-12 * np.eye(cov_in.shape[0])
        sqrt_ = np.linalg.cholesky(cov_in_)
        if update_cov:
            return mean_, cov_in_, sqrt_, ess
        return mean_, cov_in, cov_in_sqrt, ess
    except np.linalg.LinAlgError:
        return mean, cov_in, cov_in_sqrt, ws.shape[0]


def iid_weighted_update(mean_fn, log_w, samples):
    """WhiteNoiseIid.weighted_update                         ppi/policies.py:1152-1164.
    Returns (mean', std', ess, map_sequence)."""
    map_sequence = samples[np.argmax(log_w), ...]
    log_nw = log_w - logsumexp(log_w)
    nw = np.exp(log_nw)
    ess = np.exp(-logsumexp(2 * log_nw))
    corrected = samples - mean_fn[None, None, :]
    mean = np.einsum("b,bij->ij", nw, corrected)
    diff = corrected - mean[None, ...]
    std = np.sqrt(np.einsum("b,bij->ij", nw, diff ** 2))
    return mean, std, ess, map_sequence


# ----------------------------------------------------------------------------------------------
# a3-a8, a22-a23 sampling
# ----------------------------------------------------------------------------------------------
def clip_last(x, lower, upper):
    """Limiter.__call__                                      ppi/policies.py:73-75."""
    if lower is None:
        return x
    return np.clip(x, a_max=upper, a_min=lower)


def sample_matnorm(zs, mean_fn, mean, l_in, l_out, lower=None, upper=None, naive=True):
    """BaseKernel.__call__                                   ppi/policies.py:735-745.
    zs: (K, H, d) standard normals."""
    if naive:
        xs = mean_fn[None, None, :] + mean[None, :, :] + np.einsum(
            "bij, ki, jl->bkl", zs, l_in, l_out.T)
    else:
        xs = mean_fn[None, None, :] + mean[None, :, :] + np.matmul(l_in, np.matmul(zs, l_out.T))
    return clip_last(xs, lower, upper)


def sample_features(zs, mean_fn, mean, l_in, l_out, feat_t, dfeat_t=None, lower=None, upper=None,
                    naive=True):
    """BaseFeatures.__call__                                 ppi/policies.py:404-421.
    Returns (ys, ws)."""
    if naive:
        ws = mean[None, :, :] + np.einsum("bij, ki, jl->bkl", zs, l_in, l_out.T)
        xs = mean_fn[None, None, :] + np.einsum("bij,ki->bkj", ws, feat_t)
    else:
        ws = mean[None, :, :] + np.matmul(l_in, np.matmul(zs, l_out.T))
        xs = mean_fn[None, None, :] + np.matmul(feat_t, ws)
    if dfeat_t is not None:
        dxs = np.einsum("bij,ki->bkj", ws, dfeat_t) if naive else np.matmul(dfeat_t, ws)
        ys = np.concatenate((xs, dxs), axis=-1)
    else:
        ys = xs
    return clip_last(ys, lower, upper), ws


def sample_gaussian(zs, mu, sigma):
    """Cholesky-form Gaussian transform                      ppi/samplers.py:38-39,59-60.
    (MonteCarlo's numpy mvn uses an SVD factor; injected-draw parity is defined on this form,
    SURVEY 7.3 item 7.)"""
    return mu[None, :] + zs @ np.linalg.cholesky(sigma).T


def sample_iid(zs, mean_fn, mean, std, lower=None, upper=None):
    """WhiteNoiseIid.__call__                                ppi/policies.py:1166-1174."""
    xs = mean_fn[None, None, :] + mean[None, ...] + np.einsum("ij,bij->bij", std, zs)
    return clip_last(xs, lower, upper)


def _cols(t1, t2):
    t1 = t1[:, None] if t1.ndim == 1 else t1
    t2 = t2[:, None] if t2.ndim == 1 else t2
    return t1, t2


def gram(kind, t1, t2, sigma, ls=1.0, period=1.0):
    """Kernel Gram builders                     ppi/policies.py:923-932,935-974,1026-1037,1087-1091."""
    t1, t2 = _cols(np.asarray(t1, float), np.asarray(t2, float))
    if kind == "se":
        k = sigma * np.exp(-0.5 * np.power((t1 - t2.T) / ls, 2))
        if k.shape[0] == k.shape[1]:
            k += 1e-3 * sigma * np.eye(k.shape[0])
        return k
    if kind in ("matern12", "matern32", "matern52"):
        a = np.abs(t1 - t2.T)
        a[a == 0.0] = 1e-8
        if kind == "matern12":
            return sigma * np.exp(-a / ls)
        if kind == "matern32":
            d = np.sqrt(3) * a / ls
            return sigma * (1.0 + d) * np.exp(-d)
        d = np.sqrt(5) * a / ls
        return sigma * (1.0 + d + d ** 2 / 3) * np.exp(-d)
    if kind == "periodic":
        diff = t1 - t2.T
        diff[diff == 0.0] = -1e-8
        k = sigma * np.exp(-2 * np.sin(np.pi * np.abs(diff) / period) ** 2 / ls)
        if k.shape[0] == k.shape[1]:
            k += 1e-3 * sigma * np.eye(k.shape[0])
        return k
    if kind == "white":
        return sigma * ((t1.reshape(-1, 1) - t2.reshape(-1, 1).T) == 0.0)
    raise ValueError(kind)


def rff_features(t, lengthscale, order, add_bias):
    """RffFeatures.__init__/feat/dfeat                       ppi/policies.py:621-623,642-663."""
    x, w = np.polynomial.hermite.hermgauss(2 * order)
    xs = np.sqrt(2) * x[order:] / lengthscale
    ws = 2 * w[order:] / np.sqrt(np.pi)
    ph = np.einsum("p,n->np", xs, t)
    f = [np.cos(ph) * np.sqrt(ws), np.sin(ph) * np.sqrt(ws)]
    df = [-np.sin(ph) * (xs * np.sqrt(ws)), np.cos(ph) * (xs * np.sqrt(ws))]
    if add_bias:
        f.append(np.ones_like(t)[:, None])
        df.append(np.zeros_like(t)[:, None])
    return np.concatenate(f, axis=1), np.concatenate(df, axis=1)


def rbf_features(t, t_min, t_max, lengthscale, n_features, add_bias):
    """RbfFeatures.__init__/feat/dfeat                       ppi/policies.py:545-553,569-590."""
    ls = lengthscale / np.sqrt(2)
    centres = np.linspace(t_min, t_max, n_features)[:, None]
    norm = 1 / np.sqrt(np.sqrt(np.pi) * n_features * ls)
    diff = t[:, None] - centres.T
    f = [norm * np.exp(-0.5 * np.power(diff / ls, 2))]
    df = [-norm * diff / np.power(ls, 2) * np.exp(-0.5 * np.power(diff / ls, 2))]
    if add_bias:
        f.append(np.ones_like(t)[:, None])
        df.append(np.zeros_like(t)[:, None])
    return np.concatenate(f, axis=1), np.concatenate(df, axis=1)


# ----------------------------------------------------------------------------------------------
# a9 cost functions                                    optimization/functions.py:21-104
# ----------------------------------------------------------------------------------------------
def himmelblau(x):
    a = x[:, 0] * x[:, 0] + x[:, 1] - 11.0
    b = x[:, 0] + x[:, 1] * x[:, 1] - 7.0
    return -1.0 * (a * a + b * b)


def rosenbrock(x):
    return np.sum(100.0 * (x[:, 1:] - x[:, :-1] ** 2.0) ** 2 + (1.0 - x[:, :-1]) ** 2, axis=-1)


def styblinski(x):
    return 0.5 * np.sum(x ** 4.0 - 16.0 * x ** 2 + 5.0 * x, axis=-1) + 39.16599 * x.shape[1]


def rastrigin(x):
    return 10.0 * x.shape[1] + np.sum(x ** 2 - 10.0 * np.cos(2.0 * np.pi * x), axis=-1)


def noisy_sphere_matrix(dim, seed=0):
    chol = np.random.default_rng(seed).standard_normal((dim, dim))
    return chol @ chol.T


def noisy_sphere(x, a, noise):
    """noise = the (K,) standard normals the reference draws with np.random.randn (functions.py:103)."""
    return np.einsum("bi, ij, bj->b", x, a, x) + 0.01 * noise


def tracking_target(t, dim_action):
    """Test env goal trajectories                            policy_search/envs.py:65-72."""
    amps = np.linspace(-1, 1, dim_action)
    return np.concatenate(
        [amps[k] * np.sin(4 * (k + 1) * t)[:, None] / 2 for k in range(dim_action)], axis=1)


def synthetic_return(xs, target):
    """Test env tracking cost                                policy_search/envs.py:83-87.
    Uses the first target.shape[1] channels of xs (the reference's map_actions_to_joints split)."""
    q = xs[..., : target.shape[1]]
    return np.einsum("bij->b", np.power(q - target[None, :, :], 2)) / (q.shape[1] * q.shape[2])


# ------------------------------------------------------------------------------------------------
# Noise priors of SURVEY 8(f) row 2
# ------------------------------------------------------------------------------------------------
def convolve(signal, kernel):
    """ppi/policies.py:1257-1267: causal filter that reads its own already-filtered output."""
    n, h, d = signal.shape
    w_ = kernel.shape[0] - 1
    smoothed = signal.copy()
    if h > w_:
        for t in range(w_, h):
            smoothed[:, t, :] = np.einsum("bij,i->bj", smoothed[:, t - w_:t + 1, :], kernel)
    return smoothed


def sample_smooth_exploration(z, mean_fn, mean, std, beta, lower=None, upper=None):
    """SmoothExplorationNoise.__call__                       ppi/policies.py:1292-1303."""
    xs = mean_fn[None, None, :] + mean[None, ...] + std[None, ...] * convolve(z, np.array([1 - beta, beta]))
    return xs if lower is None else np.clip(xs, lower, upper)


def sample_smooth_action(z, mean_fn, mean, std, beta, lower=None, upper=None):
    """SmoothActionNoise.__call__                            ppi/policies.py:1309-1319."""
    xs = convolve(mean_fn[None, None, :] + mean[None, ...] + std[None, ...] * z, np.array([1 - beta, beta]))
    return xs if lower is None else np.clip(xs, lower, upper)


def iid_weighted_update(mean_fn, log_w, samples):
    """WhiteNoiseIid.weighted_update                         ppi/policies.py:1152-1164."""
    log_nw = log_w - logsumexp(log_w)
    nw = np.exp(log_nw)
    ess = np.exp(-logsumexp(2 * log_nw))
    mean = np.einsum("b,bij->ij", nw, samples - mean_fn[None, None, :])
    diff = samples - mean_fn[None, None, :] - mean[None, ...]
    std = np.sqrt(np.einsum("b,bij->ij", nw, diff ** 2))
    return mean, std, ess, samples[np.argmax(log_w)]


def lgds_gram(t, order, sigma):
    """LinearGaussianDynamicalSystemKernel.k                 ppi/policies.py:1371-1385."""
    from scipy.linalg import block_diag

    n = t.shape[0]
    a = np.eye(order)
    dt = t[1] - t[0]
    if order == 3:
        a[0, 1], a[0, 2], a[1, 2] = dt, 0.5 * dt ** 2, dt
    elif order == 2:
        a[0, 1] = dt
    q = np.zeros((order, order))
    q[-1, -1] = sigma
    a_ = np.kron(np.eye(n), np.eye(order))
    for i in range(1, n):
        a_ += np.kron(np.eye(n, k=-i), np.linalg.matrix_power(a, i))
    q_ = block_diag(*([1e-3 * np.eye(order)] + [q] * (n - 1)))
    gram = a_ @ q_ @ a_.T + block_diag(*([1e-6 * np.eye(order)] * n))
    return gram[::order, ::order]


def powerlaw_psd_gaussian(exponent, size, rng):
    """colorednoise.powerlaw_psd_gaussian, colorednoise==2.0.0 (ppi.yaml:23; not vendored, not
    installed -- restated from its published algorithm, Timmer & Koenig 1995).  Call site:
    ppi/policies.py:1241-1243."""
    size = list(size)
    samples = size[-1]
    f = np.fft.rfftfreq(samples)
    fmin = 1.0 / samples
    s_scale = f.copy()
    ix = np.sum(s_scale < fmin)
    if ix and ix < len(s_scale):
        s_scale[:ix] = s_scale[ix]
    s_scale = s_scale ** (-exponent / 2.0)
    w = s_scale[1:].copy()
    w[-1] *= (1 + (samples % 2)) / 2.0
    sigma = 2 * np.sqrt(np.sum(w ** 2)) / samples
    size[-1] = len(f)
    sr = rng.standard_normal(size) * s_scale
    si = rng.standard_normal(size) * s_scale
    if not (samples % 2):
        si[..., -1] = 0
        sr[..., -1] *= np.sqrt(2)
    si[..., 0] = 0
    sr[..., 0] *= np.sqrt(2)
    return np.fft.irfft(sr + 1j * si, n=samples, axis=-1) / sigma


# ---------------------------------------------------------------------------------------------
# the MPC time shift of the prior (mpc/mpc.py:64-65 -> policy.update_timesteps)
# ---------------------------------------------------------------------------------------------
def kernel_shift(gram_fn, t_old, t_new, prior, prior_inv, cov_in, mean, mean_fn, lower, upper, anneal, sigma,
                 eps=1e-5):
    """BaseKernel.update_timesteps (ppi/policies.py:761-791): GP regression of the posterior mean and
    of the covariance reduction `prior - cov_in` from the old window onto the new one.
    `gram_fn(t1, t2)` is the kernel; prior / prior_inv are the cached Gram matrix of the old window and
    its inverse (recomputed by the caller when the window length changes, :766-768).
    Returns (mean_new, cov_new, chol_new)."""
    gain = prior_inv @ (prior - cov_in) @ prior_inv.T                                   # :770-774
    cross = gram_fn(t_new, t_old)                                                       # :776
    mean_new = cross @ prior_inv @ mean                                                 # :778
    shifted = mean_new + mean_fn[None, :]
    if lower is not None:
        shifted = np.clip(shifted, lower, upper)                                        # :779
    mean_new = shifted - mean_fn[None, :]                                               # :780
    cov_new = (gram_fn(t_new, t_new) - anneal * cross @ gain @ cross.T
               + eps * sigma * np.eye(t_new.shape[0]))                                  # :782-786
    return mean_new, cov_new, np.linalg.cholesky(cov_new)


def white_kernel_shift(t_old, t_new, mean, cov_in, sigma):
    """WhiteNoiseKernel.update_timesteps (ppi/policies.py:1093-1104): kept time steps carry their
    moments over, new ones start from the prior sigma I."""
    remap = 1.0 * ((t_new.reshape(-1, 1) - t_old.reshape(1, -1)) == 0.0)
    n = t_new.shape[0]
    cov = remap @ cov_in @ remap.T + (np.eye(n) - remap @ remap.T) @ (sigma * np.eye(n))
    return remap @ mean, cov, np.linalg.cholesky(cov)


def iid_shift(t_old, t_new, mean, std, sigma_out, anneal):
    """WhiteNoiseIid.update_timesteps (ppi/policies.py:1176-1188); sigma_out = sqrt(cov_out * cov_in)
    (:1130), a (d, d) matrix the initial std is `ones @ sigma_out` of."""
    remap = 1.0 * ((t_new.reshape(-1, 1) - t_old.reshape(1, -1)) == 0.0)
    n = t_new.shape[0]
    std_new = np.ones((n, sigma_out.shape[0])) @ sigma_out
    s = np.sqrt((remap @ std) ** 2 + ((np.eye(n) - remap @ remap.T) @ std_new) ** 2)
    return remap @ mean, anneal * s + (1 - anneal) * std_new


def features_anneal(cov_in, cov_init, anneal):
    """BaseFeatures.update_timesteps (ppi/policies.py:396-402)."""
    cov = anneal * cov_in + (1 - anneal) * cov_init
    return cov, np.linalg.cholesky(cov)


def kernel_predict(mean_fn, mean, cov_in, cov_out):
    """BaseKernel.predict (ppi/policies.py:850-860): (mu_y, std_y) with std_y the square root of the
    diagonal of kron(cov_out, cov_in) in the reference's (row-major reshape) element order."""
    mu = mean_fn[None, :] + mean
    return mu, np.sqrt(np.diag(np.kron(cov_out, cov_in)).reshape(mu.shape))


# ---------------------------------------------------------------------------------------------
# quasi-Monte-Carlo base normals (ppi/samplers.py:25-44)
# ---------------------------------------------------------------------------------------------
def sobol_points(sv, shift, first, n):
    """Points first .. first + n - 1 of the scrambled Sobol sequence whose engine state is (sv (d, bits),
    shift (d)) -- scipy.stats.qmc.Sobol after construction -- in closed form: the Gray-code recurrence
    of the engine (quasi ^= sv[:, lowest zero bit of the running index]) unrolled to
    x_i = 2^-bits (shift ^ XOR_{b set in i ^ (i >> 1)} sv[:, b])."""
    sv, shift = np.asarray(sv, np.uint64), np.asarray(shift, np.uint64)
    bits = sv.shape[1]
    i = np.arange(first, first + n, dtype=np.uint64)
    gray = i ^ (i >> np.uint64(1))
    q = np.tile(shift[None, :], (n, 1))
    for b in range(bits):
        on = ((gray >> np.uint64(b)) & np.uint64(1)).astype(bool)
        q[on] ^= sv[:, b][None, :]
    return q.astype(np.float64) * 2.0 ** -bits


def qmc_base_normals(points, shrinkage=0.9999):
    """QuasiMonteCarlo.__call__(standard_gaussian=True), ppi/samplers.py:33-37."""
    from scipy.special import erfinv

    return np.sqrt(2) * erfinv(2 * shrinkage * points - 1)
