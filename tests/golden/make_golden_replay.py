"""
Golden fixtures that replay the reference's own CALLERS of the hot path:

  * optimization/run_opt.py:44-64 -- `main(args)` of the UNMODIFIED runner is executed with the whole
    argparse namespace (`alg(**vars(args))`, sampler INSTANCE, `diagonal = algorithm == "Cem"`) for every
    optimizer of ppi.algorithms.__all__ except More, with the two deterministic samplers (qmc, quad); the
    traces it saves are the fixture;
  * mpc/mpc.py:58-84 -- the UNMODIFIED `Mpc` class drives `update_timesteps -> solver(env, policy, ...) ->
    predict / map_sequence[0]` over a short episode with a shrinking horizon, on a synthetic tracking env and
    injected base normals.

    python tests/golden/make_golden_replay.py      # build container only

Shims (besides the three of make_golden.py): `matplotlib` is stubbed (run_opt.py imports it at module
top, only `--plot` uses it); `QuasiMonteCarlo` is given a seeded SciPy engine (`qmc.Sobol(seed=args.seed)`:
the reference leaves the scramble to SciPy's global entropy, which no fixture can pin); the `quad` namespace
carries `n_samples` (the reference's `quad` sub-parser forgets it and `main` would raise AttributeError).
"""
import argparse
import os
import shutil
import sys
import tempfile
import types

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from make_golden import OUT, REF, InjectedNormals, f32, import_reference  # noqa: E402

ALGS = ["Ais", "Cem", "iCem", "Reps", "Lbps", "Essps", "Mppi", "MppiUpdateCovariance"]


def namespace(algorithm, sampling, seed, dimension=5, n_iter=8, n_samples=100, out_dir=None):
    """What `python run_opt.py <algorithm> Rosenbrock <sampling>` parses (run_opt.py:96-126 defaults)."""
    return argparse.Namespace(
        algorithm=algorithm, function="Rosenbrock", dimension=dimension, n_iter=n_iter, seed=seed, plot=False,
        name="", dir=out_dir, force=True, n_elites=10, alpha=0.9, base_entropy=-100, entropy_rate=0.99,
        epsilon=0.1, delta=0.5, ess_pc=0.1, sampling=sampling, n_samples=n_samples)


def tracking_target(t, d):
    amps = np.linspace(-1, 1, d)
    return np.concatenate([amps[k] * np.sin(4 * (k + 1) * t)[:, None] / 2 for k in range(d)], axis=1)


class TrackingEnv:
    """Synthetic stand-in of mpc/wrappers.py: a callable cost on (K, H, d) action sequences; the goal
    trajectory is a function of ABSOLUTE time, so the shifting window sees a moving target."""

    def __init__(self, dt, d):
        self.dt, self.d, self.t0 = dt, d, 0
        self.rewards = np.zeros((1,))

    def set_state(self, state):
        self.t0 = int(state)

    def __call__(self, actions):
        h = actions.shape[1]
        target = tracking_target(self.dt * np.arange(self.t0, self.t0 + h), self.d)
        return np.mean((actions - target[None]) ** 2, axis=(1, 2))


def main():
    alg, pol, smp, fn = import_reference()
    g = {}

    # ---- run_opt.py --------------------------------------------------------------------------------------
    mpl = types.ModuleType("matplotlib")
    mpl.pyplot = types.ModuleType("matplotlib.pyplot")
    sys.modules.setdefault("matplotlib", mpl)
    sys.modules.setdefault("matplotlib.pyplot", mpl.pyplot)
    import run_opt  # /root/reference/optimization/run_opt.py (on sys.path through import_reference)
    from scipy.stats import qmc

    class SeededQmc(smp.QuasiMonteCarlo):
        seed = 0

        def __init__(self, dimension):
            self.sampler = qmc.Sobol(d=dimension, scramble=True, seed=SeededQmc.seed)

    run_opt.QuasiMonteCarlo = SeededQmc
    tmp = tempfile.mkdtemp(prefix="ppi_replay_")
    for sampling in ("qmc", "quad"):
        for name in ALGS:
            seed = 3
            SeededQmc.seed = seed
            args = namespace(name, sampling, seed, out_dir=tmp,
                             n_samples=100 if sampling == "qmc" else 10)
            run_opt.main(args)
            res = dict(np.load(os.path.join(tmp, f"{name}_Rosenbrock_{sampling}_{seed}_.npz"), allow_pickle=True))
            for k, v in res.items():
                if k == "episodes":
                    continue
                a = np.asarray(v, dtype=np.float64)     # Reps stores alpha as shape-(1,) arrays
                g[f"opt_{sampling}_{name}_{k}"] = a.reshape(a.shape[0], -1)[:, 0] if a.size else np.zeros(0)
    shutil.rmtree(tmp, ignore_errors=True)
    g["opt_args"] = np.array([5, 8, 100, 3])   # dimension, n_iter, n_samples (qmc), seed

    # ---- mpc/mpc.py ------------------------------------------------------------------------------------------
    sys.path.insert(0, os.path.join(REF, "mpc"))
    import importlib.util

    spec = importlib.util.spec_from_file_location("ref_mpc", os.path.join(REF, "mpc", "mpc.py"))
    ref_mpc = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref_mpc)
    rng = np.random.default_rng(77)
    d, dt, timesteps, horizon, K, n_iters, anneal = 3, 0.05, 9, 6, 64, 2, 0.5
    mean_fn = np.array([0.05, -0.1, 0.0])
    cout = np.diag([0.5, 1.0, 2.0])
    lower, upper = -0.8 * np.ones(d), 0.9 * np.ones(d)
    g.update(mpc_dt=dt, mpc_dims=np.array([d, timesteps, horizon, K, n_iters]), mpc_anneal=anneal,
             mpc_mean_fn=mean_fn, mpc_cout=cout, mpc_lower=lower, mpc_upper=upper)
    draws = []
    for ti in range(timesteps):
        h = min(horizon, timesteps - ti)
        for _ in range(n_iters):
            draws.append(f32(rng.standard_normal((K, h * d))))
    for i, z in enumerate(draws):
        g[f"mpc_z_{i}"] = z
    for use_map in (False, True):
        InjectedNormals.queue = [z.copy() for z in draws]
        policy = pol.SquaredExponentialKernel(
            time_sequence=dt * np.arange(horizon), action_dimension=d, mean=mean_fn,
            covariance_in=np.array([0.7]), covariance_out=cout, lengthscale=0.12, sampler=InjectedNormals,
            limiter=pol.Limiter(lower=lower, upper=upper))
        env = TrackingEnv(dt, d)
        mpc = ref_mpc.Mpc(env, dt, alg.Lbps(delta=0.9), policy, timesteps, horizon, K, n_iters, anneal,
                          use_map=use_map)
        actions = np.stack([mpc(ti, ti) for ti in range(timesteps)])
        tag = "map" if use_map else "mean"
        g[f"mpc_actions_{tag}"] = actions
        g[f"mpc_ess_{tag}"] = mpc.ess[:, 0].copy()
        g[f"mpc_alpha_{tag}"] = mpc.alphas[:, 0].copy()
        g[f"mpc_final_mean_{tag}"] = policy.mean.copy()
        g[f"mpc_final_cov_{tag}"] = policy.covariance_in.copy()
        g[f"mpc_prior_cond_{tag}"] = np.linalg.cond(policy.covariance_in_prior)
    np.savez_compressed(os.path.join(OUT, "reference_golden_replay.npz"), **g)
    print("wrote", len(g), "arrays; prior cond", float(g["mpc_prior_cond_mean"]))


if __name__ == "__main__":
    main()
