"""Drop-in replay tests: the construction and call sequences of the reference's own callers, restated
line by line with the ppi_b200 modules, against traces produced by the UNMODIFIED callers
(tests/golden/make_golden_replay.py):

  * optimization/run_opt.py:44-64 -- `alg(**vars(args))` with the whole CLI namespace, a sampler INSTANCE,
    `diagonal = algorithm == "Cem"`, `solver(function, policy, n_samples, n_iter)`;
  * mpc/mpc.py:58-84 -- `update_timesteps -> solver(env, policy, ...) -> predict / map_sequence[0]` over an
    episode with a shrinking horizon;
  * the policy surface `More` uses (ppi/algorithms.py:345,436): `policy.mu`, `policy.sigma`,
    `policy.smooth_update`, array-like samples / costs.

Iterated loops amplify fp32 round-off (each iteration samples from the previous fit), hence 1e-4 on the
early iterations and a looser bound on the late ones; single updates are held to 1e-5 elsewhere.
"""
import argparse
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ALGS = ["Ais", "Cem", "iCem", "Reps", "Lbps", "Essps", "Mppi", "MppiUpdateCovariance"]
EARLY_TOL = 2e-4          # first three iterations
# all eight iterations; MppiUpdateCovariance collapses the covariance onto one sample in its second iteration
# (ess = 1.0001, entropy -50 in the reference's own trace): from there on the run is round-off, in fp64 too
LATE_TOL = {"MppiUpdateCovariance": float("inf")}


def nerr(a, ref):
    a, ref = np.asarray(a, np.float64), np.asarray(ref, np.float64)
    assert a.shape == ref.shape, (a.shape, ref.shape)
    return np.max(np.abs(a - ref)) / max(np.max(np.abs(ref)), 1e-300)


@pytest.fixture(scope="module")
def g():
    return dict(np.load(os.path.join(ROOT, "tests", "golden", "reference_golden_replay.npz")))


@pytest.fixture(scope="module")
def mods():
    import torch

    assert torch.cuda.is_available()
    from ppi_b200 import algorithms, functions, policies, samplers

    return algorithms, policies, samplers, functions


def namespace(algorithm, sampling, seed, dimension=5, n_iter=8, n_samples=100):
    """What `python run_opt.py <algorithm> Rosenbrock <sampling>` parses (run_opt.py:96-126)."""
    return argparse.Namespace(
        algorithm=algorithm, function="Rosenbrock", dimension=dimension, n_iter=n_iter, seed=seed, plot=False,
        name="", dir=None, force=True, n_elites=10, alpha=0.9, base_entropy=-100, entropy_rate=0.99,
        epsilon=0.1, delta=0.5, ess_pc=0.1, sampling=sampling, n_samples=n_samples)


def run_opt_main(mods, args):
    """optimization/run_opt.py:44-64 with `algorithms`, `functions`, the policy and the samplers taken
    from ppi_b200 (the diff INTEGRATION.md shows)."""
    algorithms, policies, samplers, functions = mods
    np.random.seed(args.seed)
    alg = getattr(algorithms, args.algorithm)
    func = getattr(functions, args.function)
    solver = alg(**vars(args))

    function = func(args.dimension)
    if args.sampling == "mc":
        sampler = samplers.MonteCarlo(args.dimension)
    elif args.sampling == "qmc":
        sampler = samplers.QuasiMonteCarlo(args.dimension, seed=args.seed)   # the fixture's seeded engine
    elif args.sampling == "quad":
        sampler = samplers.CubatureQuadrature(args.dimension)

    policy = policies.GaussianPolicy(
        mu=np.ones((args.dimension,)),
        sigma=0.5 * np.eye(args.dimension),
        sampler=sampler,
        diagonal=args.algorithm == "Cem",
    )
    res = solver(function, policy, args.n_samples, args.n_iter, use_tqdm=False)
    return res, policy


def _trace(v):
    a = np.asarray([np.asarray(x, np.float64).reshape(-1)[0] for x in v], np.float64)
    return a


@pytest.mark.parametrize("sampling", ["qmc", "quad"])
@pytest.mark.parametrize("name", ALGS)
def test_run_opt_replay_deterministic_samplers(mods, g, record, name, sampling):
    dim, n_iter, n_samples, seed = (int(v) for v in g["opt_args"])
    args = namespace(name, sampling, seed, dim, n_iter, n_samples if sampling == "qmc" else 10)
    res, policy = run_opt_main(mods, args)
    keys = sorted(k[len(f"opt_{sampling}_{name}_"):] for k in g if k.startswith(f"opt_{sampling}_{name}_"))
    assert sorted(res.keys()) == keys, (sorted(res.keys()), keys)      # same telemetry as the reference's run
    worst = {}
    for k in keys:
        ref = g[f"opt_{sampling}_{name}_{k}"]
        got = _trace(res[k])
        assert got.shape == ref.shape, (k, got.shape, ref.shape)
        if not ref.size:
            continue
        fin = np.isfinite(ref)
        assert np.array_equal(np.isfinite(got), fin), k
        if name == "Reps" and k != "ent":
            # the device solver returns the exact KL = epsilon root where L-BFGS-B stops short:
            # the loop follows a slightly different temperature path; the first entropy and the trend are checked
            continue
        n_strict = 3 if k != "ent" else 4        # "ent" has the initial entropy in front
        if name == "MppiUpdateCovariance":
            # the second update collapses the covariance onto one sample (ess = 1.0001, entropy -50 in the
            # reference's own trace): from there on every determinant is round-off, in fp64 as in fp32
            n_strict = 1 if k != "ent" else 2
            if sampling == "quad":
                # ten sigma points: the collapse happens in the FIRST update already (reference trace:
                # ess[0] = 1.0, entropy -347, KL 3e91) -- only the prior entropy and the KL-free traces remain
                if k == "kl":
                    continue
                n_strict = 1
            ref, got, fin = ref[:n_strict], got[:n_strict], fin[:n_strict]
        if k == "alpha" and np.max(np.abs(ref[fin])) < 1e-6:
            # Essps with n_elites = K (the `quad` run: 10 sigma points, 10 elites): ESS(alpha) = K only at
            # alpha = 0, the reference's Brent returns round-off around it
            assert np.max(np.abs(got)) < 1e-5, got
            continue
        # KL is compared like everywhere else (relative to max(1, |kl|)); a trace that is identically ~0 in the
        # reference (Essps with n_elites = K: nothing moves) is compared absolutely
        scale = max(np.max(np.abs(ref[fin])), 1.0 if k == "kl" else 1e-6)
        early = np.max(np.abs(got[:n_strict] - ref[:n_strict])[fin[:n_strict]], initial=0.0) / scale
        late = np.max(np.abs(got - ref)[fin], initial=0.0) / scale
        worst[k] = (early, late)
    record(f"run_opt_replay[{name}-{sampling}]", **{f"{k}_first3": a for k, (a, b) in worst.items()},
           **{f"{k}_all8": b for k, (a, b) in worst.items()})
    for k, (early, late) in worst.items():
        assert early < EARLY_TOL, (name, sampling, k, early)
        assert late < LATE_TOL.get(name, 1e-2), (name, sampling, k, late)
    if name == "Reps":
        assert abs(_trace(res["ent"])[0] - g[f"opt_{sampling}_{name}_ent"][0]) < 1e-9
        assert _trace(res["mean"])[-1] < _trace(res["mean"])[0]


@pytest.mark.parametrize("name", ALGS)
def test_run_opt_replay_monte_carlo(mods, g, name):
    """`mc`: the device Philox stream replaces NumPy's global RNG, so the draws differ from the reference's;
    the run must produce the same telemetry keys and lengths, finite numbers, and make progress."""
    dim, n_iter, n_samples, seed = (int(v) for v in g["opt_args"])
    res, policy = run_opt_main(mods, namespace(name, "mc", seed, dim, 12, 256))
    keys = sorted(k[len(f"opt_qmc_{name}_"):] for k in g if k.startswith(f"opt_qmc_{name}_"))
    assert sorted(res.keys()) == keys
    for k in keys:
        got = _trace(res[k]) if len(res[k]) else np.zeros(0)
        n_ref = len(g[f"opt_qmc_{name}_{k}"])          # n_iter, n_iter + 1 ("ent": initial entropy first) or 0
        want = 12 + (n_ref - n_iter) if n_ref else 0
        assert got.shape == (want,), (k, got.shape, want)
        assert np.isfinite(got).all(), k
    m = _trace(res["mean"])
    assert m[-1] < m[0] or name == "MppiUpdateCovariance"   # the expected cost under the search distribution went down
    # same seed -> same draws (np.random.seed(args.seed) seeds the Philox stream, run_opt.py:44): the first
    # iteration is bit-identical; later ones agree to the summation order of the fp64 atomics in the SIMT
    # moment kernels, amplified by the iteration
    res2, policy2 = run_opt_main(mods, namespace(name, "mc", seed, dim, 12, 256))
    m2 = _trace(res2["mean"])
    assert m2[0] == m[0]
    rel = np.abs(m2 - m) / np.maximum(np.abs(m), 1e-300)
    print("MC_REPRO", name, [float("%.2g" % v) for v in rel])
    if name != "MppiUpdateCovariance":      # (collapses onto one sample: round-off dynamics after that, see above)
        # measured over repeated runs: <= 3e-7 on the first three iterations, Essps 2e-6 .. 3e-6 on the fourth
        # (its root-finder amplifies the atomics' summation order), <= 1e-3 after twelve
        assert np.allclose(m2[:3], m[:3], rtol=1e-5) and np.allclose(m2, m, rtol=5e-2)


class TrackingEnv:
    """The synthetic env of the fixture: cost of (K, H, d) action sequences against a goal trajectory that
    is a function of absolute time; evaluated on the device when handed a device handle."""

    device_cost = True

    def __init__(self, dt, d, fn):
        self.dt, self.d, self.t0, self.fn = dt, d, 0, fn
        self.rewards = np.zeros((1,))
        self._cache = {}

    def set_state(self, state):
        self.t0 = int(state)

    def __call__(self, actions):
        h = actions.shape[1]
        key = (self.t0, h)
        if key not in self._cache:
            self._cache[key] = self.fn.TrackingCost(self.dt * np.arange(self.t0, self.t0 + h), self.d)
        return self._cache[key](actions)


class Mpc(object):
    """mpc/mpc.py:19-84 restated (the reference file imports NumPy only; it cannot travel to the GPU box)."""

    def __init__(self, env, dt, solver, policy, timesteps, horizon, n_samples, n_iters, anneal, use_map=False):
        self.env, self.dt, self.solver, self.policy = env, dt, solver, policy
        self.timesteps, self.horizon, self.n_samples, self.n_iters = timesteps, horizon, n_samples, n_iters
        self.anneal, self.use_map = anneal, use_map
        self.ess = np.nan * np.ones((timesteps, 1))
        self.alphas = np.nan * np.ones((timesteps, 1))
        self.policy.compute_prior(self.time_sequence(0))

    def time_sequence(self, time_index):
        last_index = min(time_index + self.horizon, self.timesteps)
        return self.dt * np.arange(time_index, last_index)

    def optimize(self, state, time_index, n_iters, use_tqdm):
        t = self.time_sequence(time_index)
        self.policy.update_timesteps(t, self.anneal)
        self.env.set_state(state)
        return self.solver(self.env, policy=self.policy, n_samples=self.n_samples, n_iters=n_iters,
                           use_tqdm=use_tqdm)

    def __call__(self, state, time_index):
        res = self.optimize(state, time_index, self.n_iters, use_tqdm=False)
        if "ess" in res:
            self.ess[time_index, 0] = res["ess"][-1]
        if "alpha" in res:
            self.alphas[time_index, 0] = res["alpha"][-1]
        if self.use_map:
            return self.policy.map_sequence[0, :]
        return self.policy.predict(only_mean=True)[0, :]


@pytest.mark.parametrize("use_map", [False, True])
def test_mpc_loop_replay(mods, g, record, use_map):
    alg, pol, smp, fn = mods
    d, timesteps, horizon, K, n_iters = (int(v) for v in g["mpc_dims"])
    dt, anneal = float(g["mpc_dt"]), float(g["mpc_anneal"])
    n_draws = sum(n_iters for _ in range(timesteps))
    smp.InjectedNormals.queue = [g[f"mpc_z_{i}"].copy() for i in range(n_draws)]
    policy = pol.SquaredExponentialKernel(
        time_sequence=dt * np.arange(horizon), action_dimension=d, mean=g["mpc_mean_fn"],
        covariance_in=np.array([0.7]), covariance_out=g["mpc_cout"], lengthscale=0.12,
        sampler=smp.InjectedNormals, limiter=pol.Limiter(lower=g["mpc_lower"], upper=g["mpc_upper"]))
    env = TrackingEnv(dt, d, fn)
    mpc = Mpc(env, dt, alg.Lbps(delta=0.9), policy, timesteps, horizon, K, n_iters, anneal, use_map=use_map)
    actions = np.stack([mpc(ti, ti) for ti in range(timesteps)])
    tag = "map" if use_map else "mean"
    err_a = nerr(actions, g[f"mpc_actions_{tag}"])
    err_e = nerr(mpc.ess[:, 0], g[f"mpc_ess_{tag}"])
    err_al = nerr(mpc.alphas[:, 0], g[f"mpc_alpha_{tag}"])
    err_m = nerr(policy.mean, g[f"mpc_final_mean_{tag}"])
    err_c = nerr(policy.covariance_in, g[f"mpc_final_cov_{tag}"])
    record(f"mpc_loop_replay[{tag}]", actions=err_a, ess=err_e, alpha=err_al, mean=err_m, cov=err_c)
    # 18 chained updates and 9 GP shifts: fp32 round-off compounds (single updates are held to 1e-5 in
    # tests/test_gpu_policies.py and tests/test_gpu_shift.py)
    assert err_a < 1e-3 and err_m < 1e-3 and err_c < 1e-3
    assert err_e < 1e-3 and err_al < 1e-3
    assert not smp.InjectedNormals.queue                          # every draw was consumed, in order


def test_more_policy_surface(mods):
    """`More` stays on the host and talks to the policy through `policy.mu`,
    `policy.sigma`, `policy.smooth_update(mu_, sigma_, 1.0)` and array-valued samples / costs
    (ppi/algorithms.py:331-345,431-437).  An optimizer restating exactly those calls runs on the
    ppi_b200 GaussianPolicy, small and large D, through the ppi_b200 Base.__call__ loop."""
    alg, pol, smp, fn = mods

    class QuadraticModelSearch(alg.Base):
        name = "MORE-surface"

        def reset(self, policy):
            super().reset(policy)
            self.trace["kl"] = []

        def update(self, costs, samples, policy):
            samples, costs = np.asarray(samples), np.asarray(costs)
            assert len(samples.shape) == 2
            rewards = -1.0 * costs
            rewards -= rewards.max()
            b, Q = policy.mu, policy.sigma                              # :345
            w = np.exp(5.0 * rewards / np.abs(rewards).max())
            w /= w.sum()
            mu_ = w @ samples
            sigma_ = (w[:, None] * (samples - mu_)).T @ (samples - mu_) + 1e-3 * np.eye(b.shape[0])
            self.trace["kl"].append(float(np.sum((mu_ - b) ** 2)))
            policy.smooth_update(mu_, sigma_, 1.0)                      # :436
            assert np.allclose(policy.mu, mu_) and np.allclose(policy.sigma, sigma_)
            return policy

    for D, K in ((4, 512), (192, 4096)):
        np.random.seed(1)
        policy = pol.GaussianPolicy(np.ones(D), 0.5 * np.eye(D), smp.MonteCarlo(D))
        solver = QuadraticModelSearch()
        tr = solver(fn.Rastrigin(D), policy, K, 4, use_tqdm=False)
        assert len(tr["mean"]) == 4 and np.isfinite(tr["mean"]).all() and tr["mean"][-1] < tr["mean"][0]
        # the next draw comes from the state smooth_update installed
        x = np.asarray(policy(K)[0])
        assert np.max(np.abs(x.mean(0) - policy.mu)) < 6.0 * np.sqrt(np.max(np.diag(policy.sigma)) / K)
