"""GPU parity tests of SURVEY 8(f) row 1, the MPC time shift of the prior: `update_timesteps`,
`compute_prior`, `predict` of the kernel policies (ppi/policies.py:747-791,850-860; one fp64 launch of
`ppi_kernel_shift` per control step), the WhiteNoiseKernel / WhiteNoiseIid remaps (:1093-1104,1176-1188)
and the feature-policy annealing (:396-402), against the outputs of the unmodified reference driven
through the sequence of mpc/mpc.py:64-65 (tests/golden/reference_golden_shift.npz: five shifts, three
annealing factors, a shrinking window at the end of the episode)."""
import os

import numpy as np
import pytest

from oracle import ppi_oracle as O

pytestmark = pytest.mark.gpu
TOL = 1e-5          # fp32 moments of the weighted update
TOL64 = 1e-9        # the shift itself is fp64 on well-conditioned Gram matrices (cond <= 5e3)

KINDS = {"se": ("SquaredExponentialKernel", dict(lengthscale=0.06)),
         "matern12": ("Matern12Kernel", dict(lengthscale=0.3)),
         "matern32": ("Matern32Kernel", dict(lengthscale=0.15)),
         "matern52": ("Matern52Kernel", dict(lengthscale=0.1)),
         "periodic": ("PeriodicKernel", dict(lengthscale=0.7, period=0.37))}


def nerr(a, ref):
    a, ref = np.asarray(a, np.float64), np.asarray(ref, np.float64)
    assert a.shape == ref.shape, (a.shape, ref.shape)
    return np.max(np.abs(a - ref)) / max(np.max(np.abs(ref)), 1e-300)


@pytest.fixture(scope="module")
def g():
    return np.load(os.path.join(os.path.dirname(__file__), "golden", "reference_golden_shift.npz"))


@pytest.fixture(scope="module")
def mods():
    import torch

    assert torch.cuda.is_available()
    from ppi_b200 import policies, samplers

    return policies, samplers


def windows(g):
    H, dt, total = g["t0"].shape[0], float(g["dt"]), int(g["total"])
    for i, (idx, anneal) in enumerate(g["steps"]):
        yield i + 1, dt * np.arange(int(idx), min(int(idx) + H, total)), float(anneal)


def make(mods, g, cls, **kw):
    pol, smp = mods
    d = g["mean_fn"].shape[0]
    return getattr(pol, cls)(time_sequence=g["t0"], action_dimension=d, mean=g["mean_fn"],
                             covariance_in=np.array([0.7]), covariance_out=g["cout"], sampler=smp.MonteCarlo,
                             limiter=pol.Limiter(lower=g["lower"], upper=g["upper"]), **kw)


@pytest.mark.parametrize("name", list(KINDS))
def test_kernel_shift_sequence(mods, g, name):
    cls, kw = KINDS[name]
    policy = make(mods, g, cls, **kw)
    policy.compute_prior(g["t0"])
    assert abs(np.linalg.cond(policy.covariance_in_prior) / float(g[f"{name}_cond"]) - 1.0) < 1e-6
    policy.weighted_update(g["log_w"], g[f"{name}_xs"])
    assert nerr(policy.mean, g[f"{name}_mean_0"]) < TOL
    assert nerr(policy.covariance_in, g[f"{name}_cov_0"]) < TOL
    # (the reference's trace over a (b, d, d) stack returns one value per actuator column, :729)
    assert nerr(policy.loglikelihood(g[f"{name}_xs"]), g[f"{name}_llh"]) < 1e-4
    # the whole control sequence, errors of the fp32 update included
    for i, t, anneal in windows(g):
        policy.update_timesteps(t, anneal)
        assert policy.dim_features == t.shape[0] and policy.timesteps_match(t)
        assert nerr(policy.mean, g[f"{name}_mean_{i}"]) < 5 * TOL, (name, i)
        assert nerr(policy.covariance_in, g[f"{name}_cov_{i}"]) < 5 * TOL, (name, i)
    mu_y, s_in, s_out, s_y = policy.predict()
    assert nerr(mu_y, g[f"{name}_predict_mu"]) < 5 * TOL and nerr(s_y, g[f"{name}_predict_std"]) < 5 * TOL
    assert nerr(policy.predict(only_mean=True), g[f"{name}_predict_mu"]) < 5 * TOL


@pytest.mark.parametrize("name", list(KINDS))
def test_kernel_shift_fp64(mods, g, name):
    """The shift alone, started from the reference's own moments: fp64 agreement, step by step."""
    cls, kw = KINDS[name]
    policy = make(mods, g, cls, **kw)
    policy.compute_prior(g["t0"])
    for i, t, anneal in windows(g):
        policy.mean = g[f"{name}_mean_{i - 1}"].copy()
        policy.covariance_in = g[f"{name}_cov_{i - 1}"].copy()
        policy.update_timesteps(t, anneal)
        assert nerr(policy.mean, g[f"{name}_mean_{i}"]) < TOL64, (name, i)
        assert nerr(policy.covariance_in, g[f"{name}_cov_{i}"]) < TOL64, (name, i)
        assert nerr(policy.covariance_in_sqrt, g[f"{name}_chol_{i}"]) < TOL64, (name, i)
        assert np.allclose(policy.covariance_in_sqrt, np.tril(policy.covariance_in_sqrt))


def test_kernel_shift_same_window_is_noop(mods, g):
    policy = make(mods, g, "Matern32Kernel", lengthscale=0.15)
    policy.compute_prior(g["t0"])
    mean, cov = policy.mean, policy.covariance_in
    policy.update_timesteps(g["t0"].copy(), 0.5)
    assert policy.mean is mean and policy.covariance_in is cov


def test_kernel_shift_not_positive_definite_raises(mods, g):
    """np.linalg.cholesky semantics (ppi/policies.py:789): an indefinite shifted covariance raises."""
    policy = make(mods, g, "SquaredExponentialKernel", lengthscale=0.06)
    policy.compute_prior(g["t0"])
    policy.covariance_in = policy.covariance_in - 5.0 * np.eye(policy.dim_features)   # 'prior - cov' blows up
    with pytest.raises(np.linalg.LinAlgError):
        policy.update_timesteps(g["t0"] + float(g["dt"]), 1.0)


def test_kernel_shift_c_abi_errors(mods):
    import torch
    from ppi_b200 import _lib

    z = torch.zeros(64, dtype=torch.float64, device="cuda")
    p = z.data_ptr()
    args = [0, 1.0, 1.0, 1.0, p, 2, p, 2, p, p, p, p, 1, p, None, None, 1.0, 0.0, p, p, p, p, p, None]
    bad = list(args); bad[16] = 1.5
    with pytest.raises(_lib.PpiError):
        _lib.call("ppi_kernel_shift", *bad)
    bad = list(args); bad[14] = p                      # lower without upper
    with pytest.raises(_lib.PpiError):
        _lib.call("ppi_kernel_shift", *bad)
    bad = list(args); bad[7] = 161                     # beyond the single-block factorisation
    with pytest.raises(_lib.PpiError):
        _lib.call("ppi_kernel_shift", *bad)


def test_white_kernel_shift(mods, g):
    pol, smp = mods
    d = g["mean_fn"].shape[0]
    policy = pol.WhiteNoiseKernel(time_sequence=g["t0"], action_dimension=d, mean=g["mean_fn"],
                                  covariance_in=np.array([0.7]), covariance_out=g["cout"],
                                  sampler=smp.MonteCarlo, limiter=pol.Limiter(lower=g["lower"], upper=g["upper"]))
    policy.weighted_update(g["log_w"], g["white_xs"])
    for i, t, _ in windows(g):
        policy.update_timesteps(t)
        assert nerr(policy.mean, g[f"white_mean_{i}"]) < TOL
        assert nerr(policy.covariance_in, g[f"white_cov_{i}"]) < TOL


def test_iid_shift(mods, g):
    pol, _ = mods
    d = g["mean_fn"].shape[0]
    policy = pol.WhiteNoiseIid(time_sequence=g["t0"], action_dimension=d, mean=g["mean_fn"],
                               covariance_in=np.array([1.0]), covariance_out=g["cout"], sampler=None,
                               limiter=pol.Limiter(lower=g["lower"], upper=g["upper"]))
    policy.weighted_update(g["log_w"], g["iid_xs"])
    assert nerr(policy.mean, g["iid_mean_0"]) < TOL and nerr(policy.std, g["iid_std_0"]) < TOL
    t_old = g["t0"]
    for i, t, anneal in windows(g):
        ref = O.iid_shift(t_old, t, g[f"iid_mean_{i - 1}"], g[f"iid_std_{i - 1}"],
                          np.sqrt(g["cout"] * 1.0), anneal)
        assert nerr(ref[0], g[f"iid_mean_{i}"]) < 1e-14 and nerr(ref[1], g[f"iid_std_{i}"]) < 1e-14
        policy.update_timesteps(t, anneal)
        assert nerr(policy.mean, g[f"iid_mean_{i}"]) < TOL
        assert nerr(policy.std, g[f"iid_std_{i}"]) < TOL
        t_old = t


def test_features_anneal(mods, g):
    pol, smp = mods
    d = g["mean_fn"].shape[0]
    policy = pol.RbfFeatures(time_sequence=g["t0"], action_dimension=d, mean=g["mean_fn"],
                             covariance_in=np.array([0.5]), covariance_out=g["cout"], lengthscale=np.sqrt(3e-2),
                             n_features=5, add_bias=True, use_derivatives=False, sampler=smp.MonteCarlo,
                             limiter=pol.Limiter(lower=g["lower"], upper=g["upper"]))
    policy.weighted_update(g["log_w"], g["feat_ws"])
    policy.update_timesteps(float(g["dt"]) * np.arange(1, g["t0"].shape[0] + 1), 0.6)
    assert nerr(policy.covariance_in, g["feat_cov_1"]) < TOL
    assert nerr(policy.covariance_in_sqrt, g["feat_chol_1"]) < TOL


# ---------------------------------------------------------------------------------------------
# module-level moment matchers called directly, iterated matrix-normal MLE
# ---------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def gm():
    return np.load(os.path.join(os.path.dirname(__file__), "golden", "reference_golden_mavn.npz"))


def test_m_projection_function(mods, gm):
    pol, _ = mods
    K = gm["xs"].shape[0]
    mu, sigma, ess = pol.m_projection(gm["log_w"], gm["xs"].reshape(K, -1))
    assert nerr(mu, gm["mp_mu"]) < TOL and nerr(sigma, gm["mp_sigma"]) < TOL
    assert abs(ess - float(gm["mp_ess"])) < TOL * ess
    assert np.array_equal(sigma, sigma.T)


@pytest.mark.parametrize("iters,upd", [(1, False), (1, True), (3, True), (4, False)])
def test_m_projection_mavn_iterated(mods, gm, iters, upd):
    pol, _ = mods
    mean, cin, cout, ess = pol.m_projection_mavn(gm["log_w"], gm["xs"], gm["cov_in0"], gm["cov_out0"],
                                                 iterations=iters, update_out=upd)
    tag = f"mavn_{iters}_{int(upd)}"
    assert nerr(mean, gm[tag + "_mean"]) < TOL
    assert nerr(cin, gm[tag + "_cin"]) < TOL
    assert nerr(cout, gm[tag + "_cout"]) < TOL
    assert abs(ess - float(gm[tag + "_ess"])) < TOL * ess


def test_matrix_normal_helpers(mods, gm):
    pol, _ = mods
    kl = pol.matrix_gaussian_kl(gm["kl_m1"], gm["mavn_3_1_cin"], gm["mavn_3_1_cout"], gm["kl_m2"], gm["kl_c2"],
                                gm["cov_out0"])
    assert abs(kl - float(gm["kl"])) < 1e-9 * abs(float(gm["kl"]))
    same = pol.matrix_gaussian_kl(gm["kl_m1"], gm["mavn_3_1_cin"], gm["mavn_3_1_cout"], gm["kl_m1"],
                                  gm["mavn_3_1_cin"], gm["mavn_3_1_cout"])
    assert abs(same) < 1e-9
    assert np.array_equal(pol.vec(gm["kl_m1"]), gm["vec2"])
    assert np.array_equal(pol.vec(gm["xs"][:3].astype(np.float64)), gm["vec3"])
    z = np.random.default_rng(0).standard_normal((5, 9, 2))
    k = np.array([0.3, 0.7])
    assert nerr(pol.convolve(z, k), O.convolve(z, k)) < 1e-14
