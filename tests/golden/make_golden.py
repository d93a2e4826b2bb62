"""
Generate the golden fixtures under tests/golden/ by running the UNMODIFIED reference
(/root/reference/ppi, optimization/functions.py) on seeded, injected inputs.

Run in the build container only (the reference does not travel to the GPU box):
    python tests/golden/make_golden.py
The three compatibility shims are those of SURVEY.md 8(c): stub `colorednoise`, restore `np.NINF`,
drop `bounds` from Brent `minimize_scalar` calls (pinned SciPy 1.7.3 ignored them).
"""
import os
import sys
import types

import numpy as np

REF = os.environ.get("PPI_REFERENCE", "/root/reference")
OUT = os.path.dirname(os.path.abspath(__file__))


def import_reference():
    sys.modules.setdefault("colorednoise", types.ModuleType("colorednoise"))
    if not hasattr(np, "NINF"):
        np.NINF = -np.inf
    sys.path.insert(0, REF)
    sys.path.insert(0, os.path.join(REF, "optimization"))
    import ppi.algorithms as alg
    import ppi.policies as pol
    import ppi.samplers as smp
    import functions as fn

    _ms = alg.minimize_scalar

    def minimize_scalar_no_bounds(fun, *a, **kw):
        if kw.get("method", "").lower() == "brent":
            kw.pop("bounds", None)
        return _ms(fun, *a, **kw)

    alg.minimize_scalar = minimize_scalar_no_bounds
    return alg, pol, smp, fn


class InjectedNormals:
    """Sampler obeying the BaseSampler contract (ppi/samplers.py:9-22) that replays pre-generated
    standard normals.  Class-level queue so that `policy.sampler` (which re-instantiates the class on
    every access, ppi/policies.py:323-325) keeps the stream."""

    queue = []

    def __init__(self, dimension):
        self.d = dimension

    def __call__(self, mu, sigma, n_samples, standard_gaussian=False):
        z = InjectedNormals.queue.pop(0)
        assert z.shape == (n_samples, self.d), (z.shape, n_samples, self.d)
        if standard_gaussian:
            return z.copy()
        return mu[None, :] + z @ np.linalg.cholesky(sigma).T


def f32(x):
    """Inputs are rounded to fp32 (what the device consumes) and handed to the reference as fp64."""
    return np.asarray(x, np.float32).astype(np.float64)


def main():
    alg, pol, smp, fn = import_reference()
    rng = np.random.default_rng(0)
    g = {}

    # ---- m_projection / m_projection_mavn -----------------------------------------------
    K, D = 256, 8
    x = f32(rng.standard_normal((K, D)) * 0.7 + 1.0)
    lw = f32(-3.0 * rng.random(K))
    mu, sigma, ess = pol.m_projection(lw, x)
    g.update(mp_x=x, mp_lw=lw, mp_mu=mu, mp_sigma=sigma, mp_ess=ess)

    K, H, d = 128, 6, 3
    x3 = f32(rng.standard_normal((K, H, d)))
    lw3 = f32(-5.0 * rng.random(K))
    cin = np.eye(H) * 2.0
    cout = np.diag(f32([1e-3, 2e-3, 5e-4]))
    m, ci, co, ess3 = pol.m_projection_mavn(lw3, x3, cin, cout)
    g.update(mv_x=x3, mv_lw=lw3, mv_cout=cout, mv_mu=m, mv_cin=ci, mv_ess=ess3)

    # ---- test functions ------------------------------------------------------------------
    K, D = 32, 6
    xf = f32(rng.standard_normal((K, D)) * 1.5)
    g["fn_x"] = xf
    for name in ["Himmelblau", "Rosenbrock", "Styblinski", "Rastrigin"]:
        g["fn_" + name] = getattr(fn, name)(D)(xf)
    ns = fn.NoisySphere(D, seed=3)
    np.random.seed(11)
    noise = np.random.randn(K)
    np.random.seed(11)
    g.update(fn_NoisySphere=ns(xf), fn_ns_A=ns.A, fn_ns_noise=noise)

    # ---- every optimizer, one update on GaussianPolicy + Rosenbrock(2) (config C1) --------
    K, D = 100, 2
    z = f32(rng.standard_normal((K, D)))
    mu0, sig0 = np.ones(D), 0.5 * np.eye(D)
    g.update(c1_z=z)
    specs = {
        "Ais": dict(alpha=0.9), "Mppi": dict(alpha=0.9), "MppiUpdateCovariance": dict(alpha=0.9),
        "Reps": dict(epsilon=0.1), "Lbps": dict(delta=0.5), "Essps": dict(n_elites=10),
        "Cem": dict(n_elites=10), "iCem": dict(n_elites=10),
    }
    for name, hyper in specs.items():
        InjectedNormals.queue = [z.copy()]
        policy = pol.GaussianPolicy(mu0.copy(), sig0.copy(), InjectedNormals(D), diagonal=name == "Cem")
        solver = getattr(alg, name)(**hyper)
        solver.reset(policy)
        samples, weights = policy(K)
        samples = f32(samples)  # device sees fp32 samples
        costs = f32(fn.Rosenbrock(D)(samples))
        solver.update(costs, samples, policy)
        tr = solver.trace
        g.update({
            f"c1_{name}_samples": samples, f"c1_{name}_costs": costs,
            f"c1_{name}_mu": policy.mu, f"c1_{name}_sigma": policy.sigma,
            f"c1_{name}_ess": np.float64(tr["ess"][-1]), f"c1_{name}_kl": np.float64(tr["kl"][-1]),
            f"c1_{name}_ent": np.float64(tr["ent"][-1]),
        })
        if "alpha" in tr:
            g[f"c1_{name}_alpha"] = np.float64(np.asarray(tr["alpha"][-1]).reshape(-1)[0])
        if "weight_ent" in tr:
            g[f"c1_{name}_went"] = np.float64(tr["weight_ent"][-1])

    # ---- full loop: Ais / Essps, 3 iterations, trace -------------------------------------
    for name, hyper in {"Ais": dict(alpha=0.9), "Essps": dict(n_elites=10)}.items():
        zs = [f32(rng.standard_normal((K, D))) for _ in range(3)]
        g[f"loop_{name}_z"] = np.stack(zs)
        InjectedNormals.queue = [a.copy() for a in zs]
        policy = pol.GaussianPolicy(mu0.copy(), sig0.copy(), InjectedNormals(D))
        solver = getattr(alg, name)(**hyper)
        tr = solver(fn.Rosenbrock(D), policy, K, 3, use_tqdm=False)
        for k in ("mean", "std", "ess", "kl", "ent"):
            g[f"loop_{name}_{k}"] = np.asarray(tr[k], np.float64)
        g[f"loop_{name}_mu"] = policy.mu
        g[f"loop_{name}_sigma"] = policy.sigma

    # ---- NaN filter + all-NaN ------------------------------------------------------------
    cn = f32(rng.random(16))
    cn[[2, 7, 11]] = np.nan
    wn = f32(rng.standard_normal((16, 3)))
    c_, w_ = alg.Base.filter(cn, wn)
    g.update(nan_c=cn, nan_w=wn, nan_c_out=c_, nan_w_out=w_)

    # ---- kernel policy (SE), config-C3-like small shapes ---------------------------------
    K, H, d = 64, 8, 3
    t = 0.01 * np.arange(H)
    mean, cov_in, cov_out = pol.design_moments(np.ones(d), -np.ones(d), ratio=1000.0)
    lim = pol.Limiter(lower=-np.ones(d), upper=np.ones(d))
    for ls in (0.05, 1.0):
        tag = f"se{ls}"
        zk = f32(rng.standard_normal((K, H * d)))
        InjectedNormals.queue = [zk.copy()]
        p = pol.SquaredExponentialKernel(
            time_sequence=t, action_dimension=d, mean=mean, covariance_in=cov_in,
            covariance_out=cov_out, lengthscale=ls, sampler=InjectedNormals, limiter=lim)
        g.update({f"{tag}_t": t, f"{tag}_z": zk, f"{tag}_gram": p.covariance_in.copy(),
                  f"{tag}_lin": p.covariance_in_sqrt.copy(), f"{tag}_cout": cov_out,
                  f"{tag}_lout": p.cov_out_sqrt.copy()})
        xs, _ = p(K)
        xs = f32(xs)
        tgt = np.concatenate([np.linspace(-1, 1, d)[k] * np.sin(4 * (k + 1) * t)[:, None] / 2
                              for k in range(d)], axis=1)
        costs = f32(np.einsum("bij->b", (xs - tgt[None]) ** 2) / (H * d))
        solver = alg.Lbps(delta=0.9)
        solver.reset(p)
        solver.update(costs, xs, p)
        g.update({f"{tag}_xs": xs, f"{tag}_tgt": tgt, f"{tag}_costs": costs,
                  f"{tag}_alpha": np.float64(solver.trace["alpha"][-1]),
                  f"{tag}_ess": np.float64(solver.trace["ess"][-1]),
                  f"{tag}_mean": p.mean.copy(), f"{tag}_cin": p.covariance_in.copy(),
                  f"{tag}_cin_sqrt": p.covariance_in_sqrt.copy(), f"{tag}_map": p.map_sequence.copy()})

    # Gram matrices of the other kernels
    for cls, kind in [("Matern12Kernel", "matern12"), ("Matern32Kernel", "matern32"),
                      ("Matern52Kernel", "matern52")]:
        g[f"gram_{kind}"] = getattr(pol, cls)._k(t, t, sigma=2.0, ls=0.03)
    g["gram_periodic"] = pol.PeriodicKernel._k(t, t, sigma=2.0, ls=0.5, per=0.04)
    g["gram_white"] = pol.WhiteNoiseKernel._k(t, t, 2.0).astype(np.float64)
    g["gram_se_cross"] = pol.SquaredExponentialKernel._k(t[:5], t, sigma=2.0, ls=0.03)

    # ---- feature policies (RFF / RBF), config-C4-like small shapes -----------------------
    K, T, d = 64, 20, 2
    tt = np.linspace(0, 2, T)
    for cls, tag, kw in [("RffFeatures", "rff", dict(order=3, n_features=6)),
                         ("RbfFeatures", "rbf", dict(order=3, n_features=6))]:
        p = getattr(pol, cls)(
            time_sequence=tt, action_dimension=d, mean=np.array([0.0, 1.5707]),
            covariance_in=np.array([1e2]), covariance_out=np.diag([1e-3, 1e-3]),
            lengthscale=np.sqrt(3e-2), sampler=InjectedNormals, use_derivatives=True,
            add_bias=True, **kw)
        F = p.dim_features
        p.condition(np.zeros((1,)), np.array([[0.0, 1.5707]]))
        # NB the reference leaves covariance_in_sqrt stale after condition (policies.py:449-459); kept.
        zf = f32(rng.standard_normal((K, F * d)))
        InjectedNormals.queue = [zf.copy()]
        g.update({f"{tag}_t": tt, f"{tag}_z": zf, f"{tag}_feat": p.feat(tt), f"{tag}_dfeat": p.dfeat(tt),
                  f"{tag}_mean0": p.mean.copy(), f"{tag}_cin0": p.covariance_in.copy(),
                  f"{tag}_lin0": p.covariance_in_sqrt.copy(), f"{tag}_lout": p.cov_out_sqrt.copy()})
        ys, ws = p(K)
        ys, ws = f32(ys), f32(ws)
        tgt = np.concatenate([np.linspace(-1, 1, d)[k] * np.sin(4 * (k + 1) * tt)[:, None] / 2
                              for k in range(d)], axis=1)
        costs = f32(np.einsum("bij->b", (ys[..., :d] - tgt[None]) ** 2) / (T * d))
        solver = alg.Essps(n_elites=10)
        solver.reset(p)
        solver.update(costs, ws, p)
        g.update({f"{tag}_ys": ys, f"{tag}_ws": ws, f"{tag}_tgt": tgt, f"{tag}_costs": costs,
                  f"{tag}_alpha": np.float64(solver.trace["alpha"][-1]),
                  f"{tag}_ess": np.float64(solver.trace["ess"][-1]),
                  f"{tag}_went": np.float64(solver.trace["weight_ent"][-1]),
                  f"{tag}_mean": p.mean.copy(), f"{tag}_cin": p.covariance_in.copy()})

    # ---- WhiteNoiseIid (draws with np.random.randn directly, policies.py:1167) -------------
    K, H, d = 64, 8, 3
    p = pol.WhiteNoiseIid(time_sequence=t, action_dimension=d, mean=mean, covariance_in=cov_in,
                          covariance_out=cov_out, limiter=lim)
    zi = f32(rng.standard_normal((K, H, d)))
    orig = np.random.randn
    np.random.randn = lambda *s: zi.copy()
    xs, _ = p(K)
    np.random.randn = orig
    xs = f32(xs)
    lwi = f32(-4.0 * rng.random(K))
    g.update(iid_z=zi, iid_std0=p.std.copy(), iid_xs=xs, iid_lw=lwi)
    ess, _ = p.weighted_update(lwi, xs)
    g.update(iid_mean=p.mean.copy(), iid_std=p.std.copy(), iid_ess=np.float64(ess),
             iid_map=p.map_sequence.copy())

    np.savez_compressed(os.path.join(OUT, "reference_golden.npz"), **g)
    print(f"wrote {len(g)} arrays, "
          f"{os.path.getsize(os.path.join(OUT, 'reference_golden.npz')) / 1024:.1f} KiB")


if __name__ == "__main__":
    main()
