"""Base-normal samplers with the reference's sampler protocol (ppi/samplers.py:9-22):
`Sampler(dimension)`, `sampler(mu, sigma, n_samples, standard_gaussian=False) -> (n, d)`.

On top of the protocol, native samplers expose `device_normals(n, d)` which the
ppi_b200 policies use to keep the draw on the GPU: MonteCarlo returns Philox parameters (the
normals are generated inside the sampling kernel), InjectedNormals returns a device tensor.
QuasiMonteCarlo generates its scrambled-Sobol points and the erfinv transform on the device from the
state of a SciPy engine, CubatureQuadrature its sigma points; the reference's own
sampler classes can still be passed to the policies unchanged and are consumed through the protocol.
"""
from __future__ import annotations

import numpy as np
import torch

from . import ops
from .device import f32, require_cuda

__all__ = ["QuasiMonteCarlo", "MonteCarlo", "CubatureQuadrature", "Particles", "InjectedNormals"]


class BaseSampler(object):
    def __call__(self, mu, sigma, n_samples, standard_gaussian):
        raise NotImplementedError


class MonteCarlo(BaseSampler):
    """np.random.randn / multivariate_normal replacement (ppi/samplers.py:14-22) backed by
    on-device Philox4x32-10.  The Philox seed is taken from NumPy's global RNG at construction so
    that `np.random.seed(s)` in the runners (optimization/run_opt.py:44) still fixes the run.
    The non-standard path uses the Cholesky transform mu + z L^T (numpy's mvn factors by SVD: same
    distribution, different draws -- SURVEY 7.3 item 7)."""

    def __init__(self, dimension, seed=None, full_precision=False):
        self.d = dimension
        self.seed = int(np.random.randint(0, 2 ** 31 - 1)) if seed is None else int(seed)
        self.draw_id = 0
        # full_precision: unrounded fp32 normals, materialised per draw and consumed like injected normals
        # (fp16-pair operands, three partial products); the default keeps the normals on the fp16 grid, where
        # each one is a single exact tensor-core operand
        self.full_precision = bool(full_precision)

    def device_normals(self, n_samples, d, rows=None):
        """-> (z or None, seed, draw_id): None means 'generate with Philox inside the kernel'.
        rows = (begin, end): the caller only needs those rows of the draw (K-sharding); z then holds them."""
        draw = self.draw_id
        self.draw_id += 1
        if self.full_precision:
            begin, end = rows if rows is not None else (0, n_samples)
            return ops.philox_normal(end - begin, d, self.seed, draw, row_offset=begin, full_precision=True), self.seed, draw
        return None, self.seed, draw

    shard_aware = True   # device_normals(..., rows=) returns only the requested rows

    def __call__(self, mu, sigma, n_samples, standard_gaussian=False):
        d = self.d if mu is None else np.asarray(mu).shape[0]
        z, seed, draw = self.device_normals(n_samples, d)
        if standard_gaussian:
            if z is None:
                z = ops.philox_normal(n_samples, d, seed, draw)
            return z.cpu().numpy().astype(np.float64)
        dev = require_cuda()
        s64 = torch.as_tensor(np.ascontiguousarray(sigma, np.float64), device=dev)
        if s64.shape[0] <= 160:      # np.linalg.cholesky(sigma), ppi/samplers.py:39 -- the library's own factorisations
            l64, status = ops.cholesky_small(s64)
            status = float(status)
        else:
            l64, _, scal = ops.cholesky_blocked(s64)
            status = float(scal[0].item())
        if status != 0:
            raise np.linalg.LinAlgError("Matrix is not positive definite")
        x = ops.sample_gaussian(n_samples, mu, l64.float(), z=z, seed=seed, draw_id=draw)
        return x.cpu().numpy().astype(np.float64)


class InjectedNormals(BaseSampler):
    """Replays pre-generated standard normals (the parity harness of SURVEY 8(c)).  The queue is
    class-level because the reference re-instantiates the sampler class on every access
    (ppi/policies.py:323-325)."""

    queue = []

    def __init__(self, dimension):
        self.d = dimension

    @classmethod
    def push(cls, z):
        cls.queue.append(np.asarray(z))

    def device_normals(self, n_samples, d):
        z = InjectedNormals.queue.pop(0)
        assert z.shape[0] == n_samples and z.size == n_samples * d, (z.shape, n_samples, d)
        return f32(z).reshape(n_samples, d), 0, 0

    def __call__(self, mu, sigma, n_samples, standard_gaussian=False):
        z = InjectedNormals.queue.pop(0)
        if standard_gaussian:
            return np.array(z, dtype=np.float64)
        L = np.linalg.cholesky(sigma)
        return mu[None, :] + np.asarray(z, np.float64) @ L.T


class QuasiMonteCarlo(BaseSampler):
    """Scrambled-Sobol + erfinv base normals (ppi/samplers.py:25-44), generated on the device.
    SciPy's engine is only constructed (Joe-Kuo direction numbers + the random linear scramble and
    digital shift, a (d, 30) uint32 table); the points themselves -- bit-identical to
    `engine.random()` -- and the inverse-normal transform come from `ppi_sobol_normal`, every row a pure
    function of its index, so K-sharded ranks draw their own rows.  Like the reference, each call
    consumes the next power of two >= n_samples points of the sequence."""

    def __init__(self, dimension, seed=None):
        from scipy.stats import qmc

        self.d = dimension
        self.sampler = qmc.Sobol(d=dimension, scramble=True, seed=seed)
        self.bits = int(self.sampler._sv.shape[1])
        self._state = None
        self.generated = 0

    def _device_state(self):
        if self._state is None:
            dev = require_cuda()
            sv = np.ascontiguousarray(self.sampler._sv, dtype=np.uint32).view(np.int32)
            sh = np.ascontiguousarray(self.sampler._shift, dtype=np.uint32).view(np.int32)
            self._state = (torch.from_numpy(sv.copy()).to(dev), torch.from_numpy(sh.copy()).to(dev))
        return self._state

    def _base(self, n_samples, shrinkage=0.9999, row_offset=0, n_local=None):
        """Rows [row_offset, row_offset + n_local) of this call's (n_samples, d) base-normal matrix."""
        sv, sh = self._device_state()
        n_b2 = 2 ** int(np.ceil(np.log(n_samples) / np.log(2)))
        first = self.generated
        self.generated += n_b2
        n_local = n_samples if n_local is None else n_local
        return ops.sobol_normal(n_local, self.d, sv, sh, self.bits, first + row_offset, shrinkage)

    def device_normals(self, n_samples, d):
        assert d == self.d
        return self._base(n_samples), 0, 0

    def __call__(self, mu, sigma, n_samples, shrinkage=0.9999, standard_gaussian=False):
        base = self._base(n_samples, shrinkage).cpu().numpy().astype(np.float64)
        if standard_gaussian:
            return base
        return mu[None, :] + base @ np.linalg.cholesky(sigma).T

    @staticmethod
    def covariance_scale(n):
        return n - 1


class CubatureQuadrature(BaseSampler):
    """2d sigma points (ppi/samplers.py:47-64); ignores n_samples like the reference."""

    def __init__(self, dimension):
        self.d = dimension
        self._points = None

    @property
    def base_points(self):
        return np.sqrt(self.d) * np.concatenate((np.eye(self.d), -np.eye(self.d)))

    def device_normals(self, n_samples, d):
        if self._points is None:
            self._points = ops.cubature_points(self.d)
        return self._points, 0, 0

    def __call__(self, mu, sigma, n_samples, standard_gaussian=False):
        if standard_gaussian:
            return self.base_points
        return mu[None, :] + self.base_points @ np.linalg.cholesky(sigma).T

    @property
    def n_samples(self):
        return 2 * self.d


class Particles(object):
    """Sample re-use for iCEM (ppi/samplers.py:67-102).  Protocol-only implementation: the
    reference's add_particles unpacks a 3-D shape from the 2-D base-normal matrix and therefore
    raises as soon as particles exist; the same happens here."""

    has_particles = False

    def __init__(self, dimension):
        self.d = dimension
        self._particles = None
        self._mc = MonteCarlo(dimension)

    @property
    def particles(self):
        return self._particles

    @particles.setter
    def particles(self, particles):
        self._particles = particles
        self.has_particles = True

    def __call__(self, mu, sigma, n_samples, standard_gaussian=False):
        samples = self._mc(mu, sigma, n_samples, standard_gaussian)
        return self.add_particles(samples)

    def add_particles(self, samples):
        if self._particles is not None:
            d_n, d_t, _ = samples.shape
            n = min(self._particles.shape[0], d_n)
            samples[:n, ...] = np.asarray(self._particles)[:n, :d_t, :]
        return samples

    @staticmethod
    def covariance_scale(n):
        return n - 1
