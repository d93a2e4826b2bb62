"""Posterior policy iteration optimizers with the reference's protocol (ppi/algorithms.py):
`Alg(**hyper)`, `alg(f, policy, n_samples, n_iters, callback, use_tqdm) -> trace`,
`alg.update(costs, samples, policy) -> policy`, `alg.trace`, `alg.name`.

The weights of every method are computed on the device by one cooperative kernel
(ppi_solve_temperature / ppi_cem_weights): NaN mask, normalisation, temperature solve, normalised
weights, ESS, entropy, argmax.  The NaN filter of ppi/algorithms.py:57-68 does not compact the
arrays: failed rollouts get weight zero and are excluded from min / max / mean / std / sort, which
gives the same numbers without moving K rows.
"""
from __future__ import annotations

from warnings import warn

import numpy as np
import torch

from . import _lib, ops, samplers
from .device import DeviceArray, LazyStats, all_gather_rows, dist_info, f32, gather_global_rows, scope
from .policies import DeviceWeights, _row, _samples

__all__ = [
    "Ais",
    "Cem",
    "iCem",
    "Reps",
    "Lbps",
    "More",
    "Essps",
    "Mppi",
    "MppiUpdateCovariance",
]

EPS = np.finfo(np.float64).tiny
alpha_lower = 1e-5
alpha_upper = 5e2


def null_callback(iteration, f, actions, costs, policy):
    """Used to no-op during step function"""
    return False


def _global_costs(costs, samples):
    """-> (costs of ALL K samples on every rank as an fp32 device tensor, fp64 offset).  K-sharded device
    costs are all-gathered (4K bytes) so that each rank solves the temperature identically (SURVEY 8(e)
    step 2).  Host (fp64) costs are centred on their smallest finite value before the fp32 cast, so that a
    large common offset (1e6 +- 1) does not eat the fp32 mantissa; every weight the solver returns is
    invariant under that shift and the offset is added back to the min / max / mean statistics."""
    if isinstance(costs, DeviceArray):
        c = costs.tensor.reshape(-1)
        if dist_info()[1] > 1 and c.shape[0] != costs.global_rows:
            with scope("allgather_costs"):
                c = all_gather_rows(c, costs.global_rows)
        return c, 0.0
    c64 = np.asarray(costs, dtype=np.float64).reshape(-1)
    finite = c64[np.isfinite(c64)]
    offset = float(finite.min()) if finite.size else 0.0
    c = f32(c64 - offset).reshape(-1)
    if isinstance(samples, DeviceArray) and dist_info()[1] > 1 and c.shape[0] != samples.global_rows:
        # every rank must subtract the same constant: agree on the global minimum first
        import torch.distributed as dist

        off = torch.tensor([offset], dtype=torch.float64, device=c.device)
        dist.all_reduce(off, op=dist.ReduceOp.MIN)
        c = f32(c64 - float(off.item())).reshape(-1)
        offset = float(off.item())
        c = all_gather_rows(c, samples.global_rows)
    return c, offset


class Base(object):
    """Base class for optimizers."""

    trace = {}  # for optimization telemetry

    @classmethod
    def reset(cls, policy):
        cls.trace = {"mean": [], "std": [], "ent": []}

    @staticmethod
    def filter(costs, weights):
        """Host-array version of the NaN filter, kept for callers that use it directly
        (ppi/algorithms.py:57-68).  The update path masks on the device instead."""
        costs = np.asarray(costs)
        bad_idx = np.isnan(costs)
        if bad_idx.all():
            warn("All trajectories entered absorbing state. Not filtering.")
            return np.zeros_like(costs), weights
        good_idx = ~bad_idx
        return costs[good_idx], np.asarray(weights)[good_idx, ...]

    def __call__(self, f, policy, n_samples, n_iters, callback=null_callback, use_tqdm=True):
        """Evaluate function f with policy samples (ppi/algorithms.py:70-94)."""
        self.reset(policy)
        iterations = range(n_iters)
        if use_tqdm:
            from tqdm import tqdm as _tqdm

            iterations = _tqdm(iterations)
        self.trace["ent"] += [policy.entropy]
        for i in iterations:
            actions, weights = policy(n_samples)
            if isinstance(actions, DeviceArray) and not getattr(f, "device_cost", False):
                costs = f(np.asarray(actions))  # host callable (e.g. a MuJoCo rollout pool)
            else:
                costs = f(actions)
            self._stats = None
            policy = self.update(costs, weights, policy)
            st = self._stats
            if st is not None:  # filtered mean / std came back with the weights
                if st.all_nan:
                    warn("All trajectories entered absorbing state. Not filtering.")
                self.trace["mean"].append(st.cost_mean)
                self.trace["std"].append(st.cost_std)
            else:
                c = np.asarray(costs, dtype=np.float64)
                c = c[~np.isnan(c)] if not np.isnan(c).all() else np.zeros_like(c)
                self.trace["mean"].append(c.mean())
                self.trace["std"].append(c.std())
            if callback is not null_callback:
                # the reference hands the callback the NaN-FILTERED costs (ppi/algorithms.py:87-91)
                c = np.asarray(costs, dtype=np.float64).reshape(-1)
                bad = np.isnan(c)
                costs = np.zeros_like(c) if bad.all() else c[~bad]
            flag = callback(i, f, actions, costs, policy)
            if flag:
                break
        return self.trace

    def update(self, costs, samples, policy):
        raise NotImplementedError

    # -- device weights ---------------------------------------------------------------------------
    def _solve(self, costs, samples, mode, hyper):
        c, offset = _global_costs(costs, samples)
        # the statistics stay on the device until somebody reads one: a kernel / feature policy takes
        # them home together with its fitted moments (one synchronisation per update instead of two)
        nw, log_w, st = ops.solve_temperature(c, mode, hyper, sync=False)
        self._stats = st = LazyStats(st, offset)
        return DeviceWeights(nw, log_w, st)


class _CemBase(Base):
    def reset(self, policy):
        super().reset(policy)
        self.trace["ess"] = []
        self.trace["kl"] = []
        self.trace["ent"] = []
        self.trace["weight_ent"] = []
        policy.reset_covariance()

    def _elite_update(self, costs, samples, policy):
        c, offset = _global_costs(costs, samples)
        idx, nw, st = ops.cem_weights(c, self.n_elites, sync=False)
        self._stats = st = LazyStats(st, offset)
        log_w = torch.where(nw > 0, torch.zeros_like(nw), torch.full_like(nw, -1e12))
        w = DeviceWeights(nw, log_w, st)
        ess, kl = policy.weighted_update(w, samples)
        x, row_offset, n_total, replicated = _samples(samples)
        policy.map_sequence = _row(x, row_offset, n_total, replicated, int(st.argmax))
        self.trace["ess"] += [ess]
        self.trace["kl"] += [kl]
        self.trace["ent"] += [policy.entropy]
        self.trace["weight_ent"] += [st.weight_entropy]
        return idx, x, row_offset, n_total, replicated


class Cem(_CemBase):
    """Cross-entropy method (ppi/algorithms.py:101-128)."""

    name = "CEM"

    def __init__(self, n_elites: int, *args, **kwargs):
        self.n_elites = n_elites

    def update(self, costs, samples, policy):
        self._elite_update(costs, samples, policy)
        return policy


class iCem(_CemBase):
    """'Improved' cross-entropy method (ppi/algorithms.py:131-165)."""

    name = "iCEM"

    def __init__(self, n_elites: int, sample_reuse_pc=0.33, *args, **kwargs):
        self.n_elites = n_elites
        self.sample_reuse_pc = sample_reuse_pc
        self.n_reuse = int(sample_reuse_pc * n_elites)

    def update(self, costs, samples, policy):
        idx, x, row_offset, n_total, replicated = self._elite_update(costs, samples, policy)
        if isinstance(policy.sampler, samplers.Particles):
            # idx holds GLOBAL sample numbers: under K-sharding the re-used elites are fetched from their
            # owner ranks (every rank ends up with the same particles, as on one GPU)
            reuse = idx[: self.n_reuse]
            rows = x[reuse] if replicated else gather_global_rows(x, reuse, row_offset, n_total)
            policy.sampler.particles = rows.cpu().numpy().astype(np.float64)
            assert policy.sampler.has_particles
        return policy


class Reps(Base):
    """Relative entropy policy search (ppi/algorithms.py:168-221).  The device solver finds the
    exact stationary point of the dual, KL(nw || uniform) = epsilon, inside the reference's bounds
    (1e-5, 5e2); SciPy's L-BFGS-B stops 0.5-4 % short of it."""

    name = "REPS"

    def __init__(self, epsilon: float, *args, **kwargs):
        self.epsilon = epsilon

    def reset(self, policy):
        super().reset(policy)
        self.trace["ess"] = []
        self.trace["kl"] = []
        self.trace["alpha"] = []
        self.trace["ent"] = []

    def update(self, costs, samples, policy):
        w = self._solve(costs, samples, _lib.TEMP_REPS_KL, self.epsilon)
        ess, kl = policy.weighted_update(w, samples)
        self.trace["ess"].append(ess)
        self.trace["alpha"].append(np.array([w.stats.alpha]))  # res.x is a shape-(1,) array
        self.trace["kl"].append(kl)
        self.trace["ent"].append(policy.entropy)
        return policy


class More(Base):
    """Model-based stochastic search (ppi/algorithms.py:224-452) is outside the ported path
    (SURVEY section 2); use the reference's implementation with the ppi_b200 GaussianPolicy."""

    name = "MORE"

    def __init__(self, *args, **kwargs):
        raise NotImplementedError("More is out of scope of ppi_b200 (SURVEY section 2); "
                                  "ppi.algorithms.More runs unchanged on ppi_b200.policies.GaussianPolicy")


class MppiBase(Base):
    """Model predictive path integral control (ppi/algorithms.py:455-480)."""

    name = "MPPI"
    update_covariance: bool

    def __init__(self, alpha: float, *args, **kwargs):
        self.alpha = alpha

    def reset(self, policy):
        super().reset(policy)
        self.trace["ess"] = []
        self.trace["kl"] = []
        self.trace["ent"] = []

    def update(self, costs, samples, policy):
        w = self._solve(costs, samples, _lib.TEMP_FIXED_SHIFTED, self.alpha)
        ess, kl = policy.weighted_update(w, samples, update_covariance_in=self.update_covariance)
        self.trace["ess"].append(ess)
        self.trace["kl"].append(kl)
        self.trace["ent"].append(policy.entropy)
        return policy


class Mppi(MppiBase):
    update_covariance = False


class MppiUpdateCovariance(MppiBase):
    update_covariance = True


class Ais(Base):
    """Adaptive importance sampling (ppi/algorithms.py:495-519)."""

    name = "AIS"

    def __init__(self, alpha: float, *args, **kwargs):
        self.alpha = alpha

    def reset(self, policy):
        super().reset(policy)
        self.trace["ess"] = []
        self.trace["kl"] = []
        self.trace["ent"] = []

    def log_w(self, costs):
        w = self._solve(costs, None, _lib.TEMP_FIXED_NORMALISED, self.alpha)
        return w, self.alpha

    def update(self, costs, samples, policy):
        w = self._solve(costs, samples, _lib.TEMP_FIXED_NORMALISED, self.alpha)
        ess, kl = policy.weighted_update(w, samples)
        self.trace["ess"].append(ess)
        self.trace["kl"].append(kl)
        self.trace["ent"].append(policy.entropy)
        return policy


class Lbps(Base):
    """Lower-bound policy search (ppi/algorithms.py:522-577)."""

    name = "SNISLB"

    def __init__(self, delta: float, *args, **kwargs):
        self.delta = delta
        self.max = -np.inf

    def reset(self, policy):
        super().reset(policy)
        self.trace["ess"] = []
        self.trace["alpha"] = []
        self.trace["kl"] = []
        self.trace["ent"] = []
        self.trace["max"] = []

    def log_w(self, costs):
        w = self._solve(costs, None, _lib.TEMP_LOWER_BOUND, self.delta)
        return w, w.stats.alpha

    def update(self, costs, samples, policy):
        w = self._solve(costs, samples, _lib.TEMP_LOWER_BOUND, self.delta)
        ess, kl = policy.weighted_update(w, samples)
        self.trace["ess"].append(ess)
        self.trace["kl"].append(kl)
        self.trace["ent"].append(policy.entropy)
        self.trace["alpha"].append(w.stats.alpha)
        return policy


class Essps(Base):
    """Effective sample size policy search (ppi/algorithms.py:580-623)."""

    name = "ESSPS"

    def __init__(self, n_elites: int, *args, **kwargs):
        self.ess = int(n_elites)

    def reset(self, policy):
        super().reset(policy)
        self.trace["ess"] = []
        self.trace["kl"] = []
        self.trace["ent"] = []
        self.trace["weight_ent"] = []
        self.trace["alpha"] = []

    def update(self, costs, samples, policy):
        w = self._solve(costs, samples, _lib.TEMP_ESS_TARGET, self.ess)
        ess, kl = policy.weighted_update(w, samples)
        self.trace["ess"].append(ess)
        self.trace["kl"].append(kl)
        self.trace["ent"].append(policy.entropy)
        self.trace["alpha"].append(w.stats.alpha)
        self.trace["weight_ent"] += [w.stats.weight_entropy]
        return policy
