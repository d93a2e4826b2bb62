"""Batched cost evaluation on the device with the interface of optimization/functions.py:
`f = Rosenbrock(dim); costs = f(x)` with x (K, dim) -> (K,), lower is better.

Called with a DeviceArray (what the ppi_b200 policies return) the costs stay on the GPU and come
back as a DeviceArray; called with a NumPy array the result is copied back to a NumPy array.
`tracking_target` / `TrackingCost` are the synthetic trajectory return of the `Test` environment
(policy_search/envs.py:65-72,83-87) used in place of MuJoCo rollouts.
"""
from __future__ import annotations

import numpy as np

from . import ops
from .device import DeviceArray, f32

__all__ = ["Himmelblau", "Rosenbrock", "Rastrigin", "Styblinski", "NoisySphere"]


class Base(object):
    device_cost = True  # tells ppi_b200.algorithms to pass the device handle through
    fused_name = None   # separable functions: evaluated in the sampling epilogue of a deferred Gaussian draw

    def __init__(self, dim: int):
        self.dim = dim

    def _eval(self, x):
        raise NotImplementedError

    def __call__(self, x):
        if isinstance(x, DeviceArray):
            fused = getattr(x, "fused_cost", None)
            if fused is not None and self.fused_name is not None and x.shape[-1] == self.dim:
                c = fused(self.fused_name)     # sampler + cost in one pass, the samples never written
                if c is not None:
                    return DeviceArray(c, x.row_offset, x.global_rows)
            return DeviceArray(self._eval(x.tensor), x.row_offset, x.global_rows)
        return self._eval(f32(x)).cpu().numpy().astype(np.float64)


class Himmelblau(Base):
    f_opt = 0.0

    def _eval(self, x):
        return ops.cost_function("Himmelblau", x)


class Rosenbrock(Base):
    f_opt = 0.0

    @property
    def x_opt(self):  # optimization/functions.py:35-37 (kept as in the reference; true optimum is ones)
        return np.zeros((self.dim,))

    def _eval(self, x):
        return ops.cost_function("Rosenbrock", x)


class Styblinski(Base):
    fused_name = "Styblinski"

    @property
    def x_opt(self):
        return -2.903534 * np.ones((self.dim,))

    @property
    def f_opt(self):
        return -39.16599 * self.dim

    def _eval(self, x):
        return ops.cost_function("Styblinski", x)


class Rastrigin(Base):
    f_opt = 0.0
    A = 10.0
    fused_name = "Rastrigin"

    @property
    def x_opt(self):
        return np.zeros((self.dim,))

    def _eval(self, x):
        return ops.cost_function("Rastrigin", x)


class NoisySphere(Base):
    """x^T A x + 0.01 randn(K), A = C C^T with C ~ default_rng(seed) (optimization/functions.py:86-104).
    The evaluation noise is drawn on the device (Philox, seeded from NumPy's global RNG)."""

    sigma = 0.01
    f_opt = 0.0

    def __init__(self, dim: int, seed=0):
        super().__init__(dim)
        rng = np.random.default_rng(seed)
        chol = rng.standard_normal((dim, dim))
        self.A = chol @ chol.T
        self._A_dev = None
        self._noise_seed = None
        self._draw = 0
        self.injected_noise = None  # (K,) standard normals for parity tests

    @property
    def x_opt(self):
        return np.zeros((self.dim,))

    def _eval(self, x):
        if self._A_dev is None:
            self._A_dev = f32(self.A)
        if self.injected_noise is not None:
            noise = f32(self.injected_noise)
        else:
            if self._noise_seed is None:
                self._noise_seed = int(np.random.randint(0, 2 ** 31 - 1))
            noise = ops.philox_normal(x.shape[0], 1, self._noise_seed, self._draw).reshape(-1)
            self._draw += 1
        return ops.cost_function("NoisySphere", x, A=self._A_dev, noise=noise)


def tracking_target(t, dim_action):
    """Goal trajectories of the Test env (policy_search/envs.py:65-72)."""
    t = np.asarray(t, dtype=np.float64)
    amps = np.linspace(-1, 1, dim_action)
    return np.concatenate([amps[k] * np.sin(4 * (k + 1) * t)[:, None] / 2 for k in range(dim_action)],
                          axis=1)


class TrackingCost(object):
    """Synthetic trajectory return: mean squared tracking error on the first dt channels."""

    device_cost = True

    def __init__(self, t, dim_action):
        self.target = tracking_target(t, dim_action)
        self._target_dev = None

    def __call__(self, x):
        if self._target_dev is None:
            self._target_dev = f32(self.target)
        if isinstance(x, DeviceArray):
            return DeviceArray(ops.cost_tracking(x.tensor, self._target_dev), x.row_offset, x.global_rows)
        return ops.cost_tracking(f32(x), self._target_dev).cpu().numpy().astype(np.float64)
