"""Action priors with the reference's policy protocol (ppi/policies.py), computing on the B200.

Drop-in surface: same class names, constructor keywords (superset tolerated),
`policy(n) -> (actions, params)`, `weighted_update(log_weights, samples, update_covariance_in)
-> (ess, kl)`, readable / writable NumPy attributes (`mu`, `sigma`; `mean`, `covariance_in`,
`covariance_in_sqrt`, `covariance_out`, ...).  Sampling and moment matching run in
libppi_b200.so; the host keeps fp64 mirrors of the small state (H x H, H x d) exactly like the
reference keeps its attributes.  The host-side bookkeeping the reference does between updates
(`update_timesteps`, `condition`, `predict`) stays NumPy on those mirrors.
"""
from __future__ import annotations

import logging
import weakref

import numpy as np
import torch

from . import _lib, ops, samplers
from .device import (DeferredGaussianSamples, DeviceArray, LazyStats, all_reduce_max, all_reduce_sum, broadcast_row,
                     dist_info, f32, f64, require_cuda, scope, shard_rows, to_host64)

__all__ = [
    "RbfFeatures",
    "RffFeatures",
    "SquaredExponentialKernel",
    "WhiteNoiseKernel",
    "WhiteNoiseIid",
    "ColouredNoise",
    "SmoothActionNoise",
    "SmoothExplorationNoise",
    "Matern12Kernel",
    "Matern32Kernel",
    "Matern52Kernel",
    "PeriodicKernel",
    "LinearGaussianDynamicalSystemKernel",
    "GaussianPolicy",
    "Limiter",
    "design_moments",
]

EPS = np.finfo(np.float64).tiny  # ppi/policies.py:36
SIGMA_MIN = 1e-6                 # ppi/policies.py:37
_LOG_EPS = float(np.log(EPS))
_LOG_MAX = float(np.log(np.finfo(np.float64).max))


# ------------------------------------------------------------------------------------------------
# small host helpers (same contracts as ppi/policies.py:40-62,65-75,146-158,213-231)
# ------------------------------------------------------------------------------------------------
def design_moments(upper, lower, ratio):
    """Matrix-normal moments that explore an actuator box (ppi/policies.py:40-50)."""
    centre = 0.5 * (upper + lower)
    var_out = 0.25 * np.square(upper - lower) / ratio
    return centre, np.array([ratio]), np.diag(var_out)


def symmetric(mat):
    assert mat.ndim == 2
    return 0.5 * (mat + mat.T)


def factorized(mat):
    assert mat.ndim == 2
    return np.diag(np.diag(mat))


class Limiter(object):
    """Clip actions to the actuator range (ppi/policies.py:65-75).  The sampling kernels fuse the
    clip into their epilogue; calling the object clips a host array."""

    def __init__(self, upper, lower):
        assert upper.shape == lower.shape
        self.dim = upper.shape[0]
        self.upper, self.lower = upper, lower

    def __call__(self, x):
        assert x.shape[-1] == self.dim, f"{x.shape}, {self.dim}"
        return np.clip(x, a_max=self.upper, a_min=self.lower).reshape(x.shape)


def null_limiter(action_sequence):
    return action_sequence


def _limits(limiter):
    """(lower, upper) arrays for the kernel epilogue, or (None, None)."""
    if limiter is None or limiter is null_limiter:
        return None, None
    lo, hi = getattr(limiter, "lower", None), getattr(limiter, "upper", None)
    if lo is None or hi is None:
        raise TypeError("ppi_b200 policies need a Limiter with .lower/.upper (or null_limiter)")
    return np.asarray(lo, np.float64), np.asarray(hi, np.float64)


def _clamped_logdet(logdet, positive=True):
    """log(max(EPS, det)) of ppi/policies.py:153-154,214 evaluated from a log-determinant: keeps
    the reference's clamp at EPS and its overflow to +inf without forming the determinant."""
    if not positive or logdet < _LOG_EPS:
        return _LOG_EPS
    if logdet > _LOG_MAX:
        return float("inf")
    return float(logdet)


def multivariate_gaussian_entropy(sigma, d):
    sign, logdet = np.linalg.slogdet(sigma)
    return 0.5 * _clamped_logdet(logdet, sign > 0) + (d / 2) * (1 + np.log(2 * np.pi))


def multivariate_gaussian_kl(mu_1, sigma_1, mu_2, sigma_2):
    """KL(N1 || N2), ppi/policies.py:146-158 (host version for small D)."""
    d = sigma_1.shape[0]
    diff = mu_2 - mu_1
    s1, l1 = np.linalg.slogdet(sigma_1)
    s2, l2 = np.linalg.slogdet(sigma_2)
    return 0.5 * (_clamped_logdet(l2, s2 > 0) - _clamped_logdet(l1, s1 > 0)
                  + np.trace(np.linalg.solve(sigma_2, sigma_1)) + diff @ np.linalg.solve(sigma_2, diff)
                  - d)


def matrix_normal_entropy(covariance_in, covariance_out, d_in, d_out):
    """ppi/policies.py:219-231."""
    sf = d_out / np.trace(covariance_out)
    s_i, l_i = np.linalg.slogdet(covariance_in / sf)
    s_o, l_o = np.linalg.slogdet(sf * covariance_out)
    logdet_in = d_out * _clamped_logdet(l_i, s_i > 0)
    logdet_out = d_in * _clamped_logdet(l_o, s_o > 0)
    return 0.5 * (logdet_in + logdet_out) + (d_in * d_out / 2) * (1 + np.log(2 * np.pi))


# ------------------------------------------------------------------------------------------------
# weights handed from the optimizers to weighted_update
# ------------------------------------------------------------------------------------------------
class DeviceWeights:
    """Normalised weights of all K (global) samples on the device + the solver statistics.
    np.asarray(obj) yields the unnormalised log-weights, so foreign policies still work."""

    def __init__(self, nw, log_w, stats):
        self.nw, self.log_w, self.stats = nw, log_w, stats

    @property
    def shape(self):
        return tuple(self.nw.shape)

    def __len__(self):
        return self.nw.shape[0]

    def __array__(self, dtype=None, copy=None):
        src = self.log_w if self.log_w is not None else torch.log(self.nw)
        a = src.detach().cpu().numpy().astype(np.float64)
        return a if dtype is None else a.astype(dtype, copy=False)

    def max(self):
        return float(np.asarray(self).max())

    def min(self):
        return float(np.asarray(self).min())


_side_streams = {}


def _side_stream(dev):
    """One side stream per device for the generator of a draw that runs beside other work."""
    st = _side_streams.get(dev.index)
    if st is None:
        st = _side_streams[dev.index] = torch.cuda.Stream(device=dev)
    return st


def _weights(log_weights):
    """Anything the reference accepts as log_weights -> DeviceWeights."""
    if isinstance(log_weights, DeviceWeights):
        return log_weights
    lw = log_weights.tensor if isinstance(log_weights, DeviceArray) else log_weights
    nw, stats = ops.normalise_log_weights(lw)
    return DeviceWeights(nw, f32(lw).reshape(-1), stats)


def _samples(samples):
    """Samples handle or host array -> (local tensor, row_offset, global_rows, replicated)."""
    if isinstance(samples, DeviceArray):
        return samples.tensor, samples.row_offset, samples.global_rows, False
    t = f32(samples)
    return t, 0, t.shape[0], True  # host arrays are complete on every rank


def _local(nw, row_offset, n_local):
    return nw[row_offset:row_offset + n_local] if nw.shape[0] != n_local else nw


def _row(samples_t, row_offset, n_total, replicated, index):
    """samples[index] as a host fp64 array; under K-sharding the owner broadcasts the row."""
    if replicated or dist_info()[1] == 1:
        return samples_t[index - row_offset].detach().cpu().numpy().astype(np.float64)
    return broadcast_row(samples_t, index, n_total).cpu().numpy().astype(np.float64)


def _full_from_upper(acc):
    """Symmetric host matrix from a device accumulator that holds the upper triangle."""
    u = np.triu(to_host64(acc))
    return u + u.T - np.diag(np.diag(u))


def m_projection(log_w, samples):
    """Weighted Gaussian moment match, ppi/policies.py:78-93, on the device: weight normalisation
    (ppi_normalise_log_weights), weighted mean (ppi_wmoments_diag) and covariance (ppi_wmoments_dense).
    samples: (K, D) host array or DeviceArray.  -> (mu, sigma, ess), host fp64."""
    w = _weights(log_w)
    x, row_offset, n_total, replicated = _samples(samples)
    x = x.reshape(x.shape[0], -1)
    nw = _local(w.nw, row_offset, x.shape[0])
    s1 = ops.wmoments_diag(x, nw)
    if not replicated:
        all_reduce_sum(s1)
    cov = ops.wmoments_dense(x, nw, shift=s1.float())
    if not replicated:
        all_reduce_sum(cov)
    sigma = to_host64(cov)
    return to_host64(s1), symmetric(sigma), float(w.stats.ess)


def m_projection_mavn(log_w, samples, covariance_in, covariance_out, iterations=1, update_out=False,
                      use_tqdm=False):
    """Weighted matrix-normal moment match, ppi/policies.py:96-143, on the device, including the
    iterated MLE: `iterations` rounds, the output factor re-estimated when `update_out` (the same
    contraction kernel on the transposed samples).  -> (mean, covariance_in, covariance_out, ess)."""
    assert iterations > 0
    w = _weights(log_w)
    x, row_offset, n_total, replicated = _samples(samples)
    K_local, H, d = x.shape
    nw = _local(w.nw, row_offset, K_local)
    s1 = ops.wmoments_diag(x, nw)
    if not replicated:
        all_reduce_sum(s1)
    shift = s1.float().reshape(H, d)
    cov_out = np.asarray(covariance_out, np.float64)
    cov_in = np.asarray(covariance_in, np.float64)
    xt = shift_t = None
    for _ in range(iterations):
        acc = ops.wmoments_matnorm(x, nw, shift=shift, inv_cov_out=1.0 / np.diag(cov_out))
        if not replicated:
            all_reduce_sum(acc)
        cov_in = _full_from_upper(acc) / d
        if update_out:
            if xt is None:
                xt, shift_t = x.transpose(1, 2).contiguous(), shift.t().contiguous()
            acc_o = ops.wmoments_matnorm(xt, nw, shift=shift_t, inv_cov_out=1.0 / np.diag(cov_in))
            if not replicated:
                all_reduce_sum(acc_o)
            cov_out = _full_from_upper(acc_o) / H
    return to_host64(s1).reshape(H, d), cov_in, cov_out, float(w.stats.ess)


def vec(x):
    """Column-major vectorisation, ppi/policies.py:161-171."""
    if x.ndim == 3:
        a, b, c = x.shape
        return x.reshape((a, b * c), order="F")
    return x.reshape((-1, 1), order="F")


def matrix_gaussian_kl(mean_1, cov_in_1, cov_out_1, mean_2, cov_in_2, cov_out_2):
    """KL between two matrix normals, ppi/policies.py:174-210 (host, H x H and d x d algebra; the
    determinants are taken as clamped log-determinants so that they cannot under- or overflow)."""
    n, p = mean_1.shape
    diff = mean_2 - mean_1
    sf1, sf2 = p / np.trace(cov_out_1), p / np.trace(cov_out_2)
    cov_out_1, cov_out_2 = cov_out_1 * sf1, cov_out_2 * sf2
    cov_in_1, cov_in_2 = cov_in_1 / sf1, cov_in_2 / sf2

    def logdet(m):
        sign, ld = np.linalg.slogdet(m)
        return _clamped_logdet(ld, sign > 0)

    quad = vec(diff).T @ vec(np.linalg.solve(cov_in_2, np.linalg.solve(cov_out_2, diff.T).T))
    trace = np.trace(np.linalg.solve(cov_out_2, cov_out_1)) * np.trace(np.linalg.solve(cov_in_2, cov_in_1))
    return float(0.5 * (n * logdet(cov_out_2) - n * logdet(cov_out_1) + p * logdet(cov_in_2)
                        - p * logdet(cov_in_1) + trace + quad.item() - n * p))


def convolve(signal_samples, kernel):
    """Causal recursive filter of the smooth-noise priors over the time axis of (n, h, d) samples,
    ppi/policies.py:1257-1267: each step replaces x[t] by the kernel-weighted sum of the ALREADY
    filtered window ending at t.  Host helper; the policies apply the same filter on the device as a
    lower-triangular matrix (ema_filter_matrix)."""
    out = np.array(signal_samples, copy=True)
    width = kernel.shape[0]
    assert width > 1
    for t in range(width - 1, out.shape[1]):
        out[:, t, :] = np.tensordot(kernel, out[:, t - width + 1:t + 1, :], axes=([0], [1]))
    return out


def _draw(sampler, n_total, d):
    """Base normals for this rank's rows: (z_local or None, seed, draw_id, row_offset, n_local)."""
    begin, end = shard_rows(n_total)
    if getattr(sampler, "shard_aware", False):
        z, seed, draw = sampler.device_normals(n_total, d, rows=(begin, end))
        if z is not None:   # already this rank's rows
            return z.reshape(end - begin, d), seed, draw, begin, end - begin, n_total
    elif hasattr(sampler, "device_normals"):
        z, seed, draw = sampler.device_normals(n_total, d)
    else:  # foreign sampler obeying only the reference protocol
        z = f32(np.asarray(sampler(np.zeros((d,)), np.eye(1), n_total, standard_gaussian=True)))
        seed = draw = 0
    if z is not None:
        n_total = z.shape[0]
        begin, end = shard_rows(n_total)
        z = z.reshape(n_total, d)[begin:end].contiguous()
    return z, seed, draw, begin, end - begin, n_total


# ------------------------------------------------------------------------------------------------
# Gaussian policy (optimisation experiments)                               ppi/policies.py:234-281
# ------------------------------------------------------------------------------------------------
class GaussianPolicy(object):

    name = "Gaussian"
    small_dim = 128  # at or below: SIMT kernels + host fp64 tail; above: tcgen05 + device tail
    n_regularized = 0  # updates that took the non-PD branch of ppi/policies.py:263-268 (bench.py asserts 0)
    defer_draws = True  # Philox draws (D > small_dim) are launched when first used: cost fused into the sampler

    def __init__(self, mu, sigma, sampler, diagonal=False):
        self.sampler = sampler
        self.diagonal = diagonal  # factorized covariance
        # like the reference (ppi/policies.py:238-241) mu / sigma are kept by reference; the private copy
        # sigma_init lives in HBM for large D (one H2D of the caller's -- possibly pinned -- array, which
        # is also the first device copy of sigma) and comes back to the host only if somebody reads it
        sigma = np.asarray(sigma, dtype=np.float64)
        self._dim = sigma.shape[0]
        self._sigma_init_host = self._sigma_init_dev = None
        self._set(np.asarray(mu, dtype=np.float64), sigma)
        if self._dim > self.small_dim and torch.cuda.is_available():
            self._sigma_init_dev = self._sigma64_dev = torch.as_tensor(sigma).to(require_cuda())
        else:
            self._sigma_init_host = sigma.copy()

    @property
    def sigma_init(self):
        if self._sigma_init_host is None:
            self._sigma_init_host = to_host64(self._sigma_init_dev)
        return self._sigma_init_host

    # -- state: device fp32 copies used by the kernels, host fp64 mirrors read by callers --------
    def _set(self, mu, sigma, chol64=None, logdet=None):
        self._mu_host, self._sigma_host = mu, sigma
        self._mu_dev = self._sigma_dev = self._chol_dev = self._sigma64_dev = None
        self._chol64_dev, self._logdet = chol64, logdet
        self._version = getattr(self, "_version", 0) + 1  # invalidates base-normal stores of earlier draws

    @property
    def mu(self):
        if self._mu_host is None:
            self._mu_host = to_host64(self._mu_dev)
        return self._mu_host

    @mu.setter
    def mu(self, value):
        self._mu_host, self._mu_dev = np.asarray(value, np.float64), None
        self._version += 1

    @property
    def sigma(self):
        if self._sigma_host is None:
            self._sigma_host = to_host64(self._sigma64_d())
        return self._sigma_host

    @sigma.setter
    def sigma(self, value):
        self._sigma_host, self._sigma_dev = np.asarray(value, np.float64), None
        self._chol_dev = self._chol64_dev = self._logdet = self._sigma64_dev = None
        self._version += 1

    def _mu_d(self):
        if self._mu_dev is None:
            self._mu_dev = f32(self._mu_host)
        return self._mu_dev

    def _sigma64_d(self):
        if self._sigma64_dev is not None:   # the constructor's upload, valid until sigma changes
            return self._sigma64_dev
        if self._sigma_dev is not None:
            return self._sigma_dev.double()
        if self._sigma_host is None:  # after a whitened update only the factor exists: sigma = L L^T
            return ops.tri_gram(self._chol64_dev)
        return f64(self._sigma_host)

    def _factor(self):
        """Lower Cholesky factor of sigma on the device (fp64 factorisation, fp32 copy for the
        sampling kernels) + log-determinant; cached until sigma changes."""
        if self._chol_dev is None:
            if self._chol64_dev is None:
                s64 = self._sigma64_d()
                if s64.shape[0] <= 160:
                    l64, status = ops.cholesky_small(s64)
                    l32 = logdet = None
                else:  # blocked fp64 factorisation (dense64.cu), fp32 copy and log det in the same call
                    l64, l32, scal = ops.cholesky_blocked(s64, want32=True)
                    status, logdet = (float(v) for v in scal[:2].cpu())
                if status != 0:
                    raise np.linalg.LinAlgError("sigma is not positive definite")
                self._chol64_dev, self._chol_dev, self._logdet = l64, l32, logdet
            if self._chol_dev is None:
                self._chol_dev = self._chol64_dev.float().contiguous()
            if self._logdet is None:
                self._logdet = float(2.0 * torch.log(torch.diagonal(self._chol64_dev)).sum().item())
        return self._chol_dev

    @property
    def entropy(self):
        d = self._dim
        if d <= self.small_dim:
            return multivariate_gaussian_entropy(self.sigma, d)
        self._factor()
        return 0.5 * _clamped_logdet(self._logdet) + (d / 2) * (1 + np.log(2 * np.pi))

    def __call__(self, n_samples):
        d = self._dim
        if not hasattr(self.sampler, "device_normals"):
            # foreign sampler: it owns the transform (ppi/policies.py:248)
            x = np.asarray(self.sampler(self.mu, self.sigma, n_samples, standard_gaussian=False))
            begin, end = shard_rows(x.shape[0])
            handle = DeviceArray(f32(x[begin:end]), begin, x.shape[0])
            return handle, handle
        if (self._chol_dev is None and self._pre is None and self.defer_draws and d > self.small_dim
                and not self.diagonal and _lib.load().ppi_get_option(_lib.OPT_GEMM_SCHEME) == _lib.GEMM_FP16X2):
            # sigma is still on the host (a fresh policy, or the caller replaced it): the base normals do not
            # depend on it, so their generator starts on the side stream now and runs beside the 4 D^2-byte
            # upload (copy engine) and the factorisation that _factor() enqueues below
            self._prefetch_next(n_samples)
        z, seed, draw, begin, n_local, n_total = _draw(self.sampler, n_samples, d)
        if d > self.small_dim and not self.diagonal and _lib.load().ppi_get_option(_lib.OPT_GEMM_SCHEME) == _lib.GEMM_FP16X2:
            # keep the base normals in HBM: weighted_update on this very handle contracts z, not x
            pre, self._pre = self._pre, None
            if pre is not None and not (z is None and pre.key == (id(self.sampler), seed, draw, begin, n_local, d)):
                pre = None          # a prefetched store of another draw: not ours
            if z is None and self.defer_draws:
                # Philox draw: nothing is launched yet -- the cost function decides whether x is ever written
                slot = None if pre is None else pre.slot

                def after(base, slot=slot, n_samples=n_samples):
                    if slot is not None:
                        self._pre_owner[slot] = base              # this store now backs a launched draw
                        self._pre_pending[slot] = None
                    if self.prefetch_normals == "sample":
                        self._prefetch_next(n_samples)            # right behind the sampling GEMMs

                handle = DeferredGaussianSamples(self, self._version, self._mu_d(), self._factor(), n_local, d,
                                                 begin, n_total, seed, draw, after=after, pre=pre)
                if slot is not None:
                    self._pre_pending[slot] = weakref.ref(handle)
                return handle, handle
            if pre is not None:
                x, base = ops.sample_gaussian_pre(n_local, self._mu_d(), self._factor(), pre)
                self._pre_owner[pre.slot] = base
            else:
                x, base = ops.sample_gaussian_keep(n_local, self._mu_d(), self._factor(), z=z, seed=seed,
                                                   draw_id=draw, row_offset=begin)
            handle = DeviceArray(x, begin, n_total)
            handle.white = (self, self._version, base)
            if z is None and self.prefetch_normals == "sample":
                self._prefetch_next(n_samples)
            return handle, handle
        x = ops.sample_gaussian(n_local, self._mu_d(), self._factor(), z=z, seed=seed, draw_id=draw,
                                row_offset=begin)
        handle = DeviceArray(x, begin, n_total)
        return handle, handle

    # Base normals of the NEXT draw generated ahead of time (they depend on the sampler's (seed, draw
    # counter, row) only, not on the policy): "sample" = right behind the sampling GEMMs, on a side stream,
    # under the cost / temperature / moment kernels of this update; "update" = behind the covariance
    # kernel, under the dense fp64 tail; None = off.  Only for draws of at least prefetch_min_elements.
    # Off by default: measured at C5 it buys 0.3 ms of 22.9 ("sample") or nothing ("update") -- the
    # generator fills every SM, so the cooperative temperature solver and the cuSOLVER panels queue behind
    # it, and under the 1 kW cap concurrent kernels share one power budget.
    prefetch_normals = None
    prefetch_min_elements = 1 << 26
    _pre = None
    _pre_owner = [None, None]
    _pre_pending = [None, None]

    def _prefetch_next(self, n_samples):
        sampler = self.sampler
        d = self._dim
        begin, end = shard_rows(n_samples)
        n_local = end - begin
        if (not isinstance(sampler, samplers.MonteCarlo) or sampler.full_precision
                or n_local * d < self.prefetch_min_elements or d <= self.small_dim):
            return
        dev = require_cuda()
        st = self.__dict__.setdefault("_pre_state", {})
        if st.get("shape") != (n_local, d):
            st.update(shape=(n_local, d), stores=[None, None], side=_side_stream(dev), slot=0)
            self._pre_owner = [None, None]
            self._pre_pending = [None, None]
        slot = st["slot"] = 1 - st["slot"]
        if st["stores"][slot] is None:      # allocated on first use: a policy that prefetches once holds one store
            st["stores"][slot] = ops._aligned_bytes(_lib.query("ppi_zstore_bytes", n_local, d, 0), dev)
        owner = self._pre_owner[slot]
        if owner is not None:
            owner.valid = False     # a handle that still holds this store falls back to the x-space update
            self._pre_owner[slot] = None
        pending = self._pre_pending[slot]
        if pending is not None and pending() is not None:
            pending()._pre = None   # a deferred draw that has not run yet: it will generate its own store
        self._pre_pending[slot] = None
        key = (id(sampler), sampler.seed, sampler.draw_id, begin, n_local, d)
        self._pre = ops.philox_prefetch(n_local, d, sampler.seed, sampler.draw_id, begin, st["stores"][slot],
                                        st["side"], key)
        self._pre.slot = slot

    def weighted_update(self, log_weights, samples, update_covariance_in=True):
        w = _weights(log_weights)
        if isinstance(samples, DeferredGaussianSamples) and samples.white is None:
            samples.tensor    # never costed nor read: launch the draw now (keeps its base normals)
        white = getattr(samples, "white", None)
        if (white is not None and white[0] is self and white[1] == self._version and update_covariance_in
                and not self.diagonal and white[2].valid):
            # the policy's own draw: contract the stored base normals -- x itself is not needed (and a
            # deferred draw whose cost was fused never wrote it)
            n_local = samples.shape[0]
            nw = _local(w.nw, samples.row_offset, n_local)
            return self._weighted_update_white(w, white[2], nw, samples.global_rows)
        x, row_offset, n_total, replicated = _samples(samples)
        nw = _local(w.nw, row_offset, x.shape[0])
        d = x.shape[1]
        if d > self.small_dim:
            return self._weighted_update_large(w, x, nw, replicated, update_covariance_in)
        mu_old, sigma_old = self.mu.copy(), self.sigma.copy()

        # pass 1: weighted mean, centred on the old mean for fp32 accuracy (policies.py:88)
        shift_old = self._mu_d()
        s1 = ops.wmoments_diag(x, nw, shift1=shift_old)
        if not replicated:
            all_reduce_sum(s1)
        mu_new_dev, _ = ops.finalize_gaussian(s1=s1, shift=shift_old, want_sigma=False)
        mu_new = mu_old + s1.cpu().numpy()

        # pass 2: covariance centred on the new mean (policies.py:89-92)
        sigma_new_host = chol64 = None
        pd = True
        if not update_covariance_in:
            pass
        elif self.diagonal:
            _, s2 = ops.wmoments_diag(x, nw, shift2=mu_new_dev, second=True)
            if not replicated:
                all_reduce_sum(s2)
            var = s2.cpu().numpy()
            sigma_new_host = np.diag(var)
            pd = bool(np.all(var > 0.0) and np.all(np.isfinite(var)))
            if pd:
                chol64 = torch.diag(torch.sqrt(s2))
        else:
            acc = ops.wmoments_matnorm(x.reshape(x.shape[0], d, 1), nw, shift=mu_new_dev.reshape(d, 1))
            if not replicated:
                all_reduce_sum(acc)
            fin = ops.finalize_matnorm(torch.zeros(d, dtype=torch.float64, device=x.device), acc, d, 1)
            sigma_new_host = fin["cov64"].cpu().numpy()
            pd = int(fin["status"].item()) == 0
            chol64 = fin["chol64"]

        # accept / regularise exactly like ppi/policies.py:258-268
        self._mu_host, self._mu_dev = mu_new, mu_new_dev
        if update_covariance_in:
            if pd:
                self._sigma_host, self._sigma_dev = sigma_new_host, None
                self._chol64_dev, self._chol_dev, self._logdet = chol64, None, None
            else:
                logging.warning("Fitted multivariate normal not positive definite. Regularizing covariance.")
                self.n_regularized += 1
                self.sigma = sigma_old + SIGMA_MIN * np.eye(d)
        ess = float(w.stats.ess)
        self._version += 1
        return ess, float(multivariate_gaussian_kl(self.mu, self.sigma, mu_old, sigma_old))

    def _weighted_update_white(self, w, base, nw, n_total=None):
        """M-projection in whitened coordinates for samples this policy drew itself (x = mu + L z with
        the base normals still in HBM, ppi_sample_gaussian_keep).  With m = sum nw z and
        C = sum nw z z^T - m m^T:  mu' = mu + L m  (ppi/policies.py:88),  sigma' = L C L^T (:89-92), so
        chol(sigma') = L chol(C), log det sigma' = log det sigma + log det C, and
        KL(new || old) = (tr C + |m|^2 - D - log det C) / 2 (:146-158) -- no pass over x, no triangular
        solve.  sigma itself is formed lazily (L' L'^T) when a caller reads it."""
        d = base.dim
        self._factor()
        l_old, mu_old_dev, old_logdet = self._chol64_dev, self._mu_d(), self._logdet
        # first moments: row 0 = m = sum nw z, row 1 = diag(S) with round-to-nearest products (one tensor, so
        # that K-sharded ranks reduce both in one collective)
        m12 = torch.zeros((2, d), dtype=torch.float64, device=mu_old_dev.device)
        m, diag = m12[0], m12[1]
        if self.prefetch_normals == "moments" and n_total is not None and not base.injected:
            self._prefetch_next(n_total)        # under the moment kernels and the tail
        _, s = ops.wmoments_white(base, nw, s1=m, s2=diag)
        if self.prefetch_normals == "update" and n_total is not None and not base.injected:
            self._prefetch_next(n_total)
        if dist_info()[1] > 1:
            with scope("allreduce_moments"):
                all_reduce_sum(m12)
                all_reduce_sum(s)
        with scope("dense_tail"):
            # C = S - m m^T, G = chol(C), L' = L G, mu' = mu + L m and the three scalars of the KL in ONE call of
            # hand-written fp64 kernels (dense64.cu); one 32-byte device->host read decides the branch
            l_new64, l_new32, mu_new, scal = ops.white_finish(s, m, l_old, mu_old_dev, diag=diag)
            status, logdet_c, trace_c, msq = (float(v) for v in scal.cpu())
        if status == 0:
            self._mu_host, self._mu_dev = None, mu_new
            self._sigma_host = self._sigma_dev = self._sigma64_dev = None
            self._chol64_dev, self._chol_dev, self._logdet = l_new64, l_new32, old_logdet + logdet_c
            kl = 0.5 * (_clamped_logdet(old_logdet) - _clamped_logdet(self._logdet) + trace_c + msq - d)
        else:  # ppi/policies.py:263-268: keep the old covariance, regularised
            logging.warning("Fitted multivariate normal not positive definite. Regularizing covariance.")
            self.n_regularized += 1
            self._mu_host, self._mu_dev = None, ops.tri_matvec(l_old, m, mu_old_dev)
            sigma_old = ops.tri_gram(l_old)
            self._sigma_host, self._chol_dev, self._chol64_dev, self._logdet = None, None, None, None
            # kept in fp64: the 1e-6 regulariser is 17 ulps of an fp32 entry of size one
            self._sigma_dev = None
            sigma_old.diagonal().add_(SIGMA_MIN)
            self._sigma64_dev = sigma_old
            self._factor()
            # tr(sigma_old^-1 (sigma_old + eps I)) = D + eps |L_old^-1|_F^2
            eye = torch.eye(d, dtype=torch.float64, device=l_old.device)
            tr = d + SIGMA_MIN * float(ops.tri_solve_sumsq(l_old, eye, None)[0].item())
            kl = 0.5 * (_clamped_logdet(old_logdet) - _clamped_logdet(self._logdet) + tr + msq - d)
            # (mu' - mu)^T sigma^-1 (mu' - mu) = |m|^2
        self._version += 1
        return float(w.stats.ess), kl

    def _weighted_update_large(self, w, x, nw, replicated, update_covariance_in):
        """d > small_dim: every array stays in HBM (tcgen05 covariance, device fp64 Cholesky /
        log-det / KL); the only device -> host traffic is the two scalars returned.  `mu` /
        `sigma` are copied to the host lazily when a caller reads them."""
        d = x.shape[1]
        self._factor()
        mu_old_dev, old_chol64, old_logdet = self._mu_d(), self._chol64_dev, self._logdet
        sigma_old_host, sigma_old_dev = self._sigma_host, self._sigma_dev

        # pass 1: weighted mean, centred on the old mean for fp32 accuracy (policies.py:88)
        absdev = torch.zeros(1, dtype=torch.float32, device=x.device)
        s1 = ops.wmoments_diag(x, nw, shift1=mu_old_dev, absdev=absdev)
        if not replicated:
            all_reduce_sum(s1)
            all_reduce_max(absdev)
        mu_new_dev, _ = ops.finalize_gaussian(s1=s1, shift=mu_old_dev, want_sigma=False)
        self._mu_host, self._mu_dev = None, mu_new_dev
        # the new mean is a convex combination of the samples: |x - mu_new| <= 2 max |x - mu_old|
        dev_bound = 2.0 * absdev

        # pass 2: covariance centred on the new mean (policies.py:89-92)
        if update_covariance_in:
            if self.diagonal:
                _, s2 = ops.wmoments_diag(x, nw, shift2=mu_new_dev, second=True)
                if not replicated:
                    all_reduce_sum(s2)
                cov = torch.diag(s2).float()
                chol64 = torch.diag(torch.sqrt(s2))
                pd = bool((torch.isfinite(s2) & (s2 > 0)).all().item())
            else:
                cov = ops.wmoments_dense(x, nw, shift=mu_new_dev, dev_bound=dev_bound)
                if not replicated:
                    all_reduce_sum(cov)
                chol64, chol32_new, scal = ops.cholesky_blocked(cov.double(), want32=True)
                status, logdet_new = (float(v) for v in scal[:2].cpu())
                pd = status == 0
            # accept / regularise exactly like ppi/policies.py:258-268
            if pd:
                self._sigma_host, self._sigma_dev, self._sigma64_dev = None, cov, None
                self._chol64_dev, self._chol_dev = chol64, None
                self._logdet = float(2.0 * torch.log(torch.diagonal(chol64)).sum().item())
                if not self.diagonal:
                    self._chol_dev, self._logdet = chol32_new, logdet_new
            else:
                logging.warning("Fitted multivariate normal not positive definite. Regularizing covariance.")
                self.n_regularized += 1
                if sigma_old_host is not None:
                    self.sigma = sigma_old_host + SIGMA_MIN * np.eye(d)
                else:
                    sigma_old = self._sigma64_d()  # device state still describes the old covariance
                    self._sigma_host = self._sigma_dev = None
                    self._sigma64_dev = sigma_old + SIGMA_MIN * torch.eye(d, dtype=torch.float64, device=x.device)
                    self._chol_dev = self._chol64_dev = self._logdet = None
                self._factor()
        # KL(new || old), ppi/policies.py:146-158,269, from the two Cholesky factors:
        # tr(S_old^-1 S_new) = |L_old^-1 L_new|_F^2 ; diff^T S_old^-1 diff = |L_old^-1 diff|^2
        l_new, logdet_new = self._chol64_dev, self._logdet
        diff = (mu_old_dev - mu_new_dev).double().contiguous()
        tr_quad = float(ops.tri_solve_sumsq(old_chol64, l_new.contiguous(), diff).sum().item())
        kl = 0.5 * (_clamped_logdet(old_logdet) - _clamped_logdet(logdet_new) + tr_quad - d)
        self._version += 1
        return float(w.stats.ess), kl

    @staticmethod
    def smooth(x, y, alpha):
        return alpha * x + (1 - alpha) * y

    def smooth_update(self, mu, sigma, alpha):
        self.mu = self.smooth(mu, self.mu, alpha)
        self.sigma = self.smooth(sigma, self.sigma, alpha)

    def reset_covariance(self):
        self.sigma = self.sigma_init.copy()


# ------------------------------------------------------------------------------------------------
# Matrix-normal primitives                                                 ppi/policies.py:288-390
# ------------------------------------------------------------------------------------------------
class BasePrimitive(object):
    """Kernel- and feature-based policies: matrix-normal over (dim_features, dim_out)."""

    dim_features: int
    dim_out: int
    map_sequence: np.ndarray = None

    def __init__(self, time_sequence, action_dimension, mean, covariance_in, covariance_out, sampler,
                 limiter=null_limiter, use_derivatives=False):
        self.t = time_sequence
        self.mean = mean
        self.dim_out = action_dimension
        self.covariance_in_init = covariance_in.copy()
        self.covariance_in = covariance_in
        self.covariance_in_sqrt = self._chol(covariance_in)
        self.covariance_out = covariance_out
        self.cov_out_sqrt = self._chol(self.covariance_out)
        self._sampler = sampler
        self._sampler_cache = {}
        self.limiter = limiter
        self.use_derivatives = use_derivatives

    @staticmethod
    def _chol(a):
        """np.linalg.cholesky semantics on the device (single-block fp64 kernel)."""
        a = np.asarray(a, np.float64)
        if a.shape[0] > 160:
            return np.linalg.cholesky(a)  # beyond the single-block kernel; not on the update path
        l, status = ops.cholesky_small(a)
        if status != 0:
            raise np.linalg.LinAlgError("Matrix is not positive definite")
        return l.cpu().numpy()

    def reset_covariance(self):
        self.covariance_in = self.covariance_in_init.copy()
        self.covariance_in_sqrt = self._chol(self.covariance_in)

    @property
    def sampler(self):
        # the reference instantiates the class on every access (ppi/policies.py:323-325); the
        # instance is cached here so that the Philox draw counter advances between calls
        key = self.dim_sample
        if key not in self._sampler_cache:
            self._sampler_cache[key] = self._sampler(key)
        return self._sampler_cache[key]

    @property
    def dim_sample(self):
        return self.dim_features * self.dim_out

    @property
    def mu_z(self):
        return np.zeros((self.dim_sample,))

    @property
    def sigma_z(self):
        return np.eye(self.dim_sample)

    def condition(self, t, action):
        dim_t = t.shape[0]
        assert action.shape == (dim_t, self.dim_out), f"{action.shape} ~= {dim_t, self.dim_out}"
        self._condition(t, action)

    def _condition(self, t, action):
        raise NotImplementedError

    def update(self, mean, covariance_in):
        assert mean.shape == (self.dim_features, self.dim_out)
        assert covariance_in.shape == (self.dim_features, self.dim_features)
        self.mean = mean
        self.covariance_in = covariance_in
        self.covariance_in_sqrt = self._chol(covariance_in)

    def base_sample(self, n_samples):
        """(n, dim_features, dim_out) standard normals as a host array (ppi/policies.py:361-365)."""
        z = self.sampler(self.mu_z, self.sigma_z, n_samples, standard_gaussian=True)
        return np.asarray(z).reshape((-1, self.dim_features, self.dim_out))

    @staticmethod
    def smooth(x, y, alpha):
        return alpha * x + (1 - alpha) * y

    def smooth_update(self, mean, covariance_in, alpha):
        self.mean = self.smooth(mean, self.mean, alpha)
        self.covariance_in = self.smooth(covariance_in, self.covariance_in, alpha)

    @property
    def entropy(self):
        return 0.0  # ppi/policies.py:379-384

    def compute_prior(self, plan_time):
        self.t = plan_time

    # -- shared device pieces ---------------------------------------------------------------------
    def _l_out_is_diag(self):
        c = np.asarray(self.cov_out_sqrt)
        hit = self.__dict__.get("_diag_flag")
        if hit is None or hit[0] is not c or not np.array_equal(hit[1], c):
            hit = self._diag_flag = (c, c.copy(), bool(np.count_nonzero(c - np.diag(np.diag(c))) == 0))
        return hit[2]

    def _update_consts(self, shift1, H, d):
        """Device copies of the per-update constants (the centring shift in fp32 and fp64, 1 / diag(cov_out)),
        rebuilt only when the host arrays they come from were replaced or edited."""
        cov_out = np.asarray(self.covariance_out, np.float64)
        # the full content is the key (H x d fp64, a few KB): any in-place edit of mean_fn / cov_out is seen
        key = (None if shift1 is None else (shift1.shape, np.ascontiguousarray(shift1).tobytes()),
               np.diagonal(cov_out).tobytes(), H, d)
        c = self.__dict__.get("_consts")
        if c is None or c[0] != key:
            s32 = None if shift1 is None else f32(np.ascontiguousarray(shift1))
            s64 = None if shift1 is None else f64(np.ascontiguousarray(shift1)).reshape(-1)
            c = self._consts = (key, s32, s64, f32(1.0 / np.diag(cov_out)))
        return c[1], c[2], c[3]

    _fit_dev = None

    def _seed_fit(self, with_factor):
        """The fp32 device copies the fused update left behind are what the next draw needs: register them
        against the fresh host mirrors so that the sampler does not upload what was just downloaded."""
        from .device import seed_f32_cache

        if self._fit_dev is not None:
            mean32, chol32 = self._fit_dev
            seed_f32_cache(self.mean, mean32)
            if with_factor:
                seed_f32_cache(self.covariance_in_sqrt, chol32)
            self._fit_dev = None

    def _mavn_update(self, w, x, row_offset, replicated, shift1, jitter, extra=None):
        """m_projection_mavn (ppi/policies.py:96-143, iterations=1, update_out=False) on the
        device: returns host fp64 (mean, cov_in, chol or None, pd[, map row]).  extra="argmax": the most
        likely sample rides along in the single device->host read of the results."""
        K_local, H, d = x.shape
        nw = _local(w.nw, row_offset, K_local)
        s32, s64, inv_out = self._update_consts(shift1, H, d)
        if (replicated or dist_info()[1] == 1) and H <= 128:
            # one rank: the whole match is one C call and one device->host read; solver statistics that
            # are still on the device come home in the same read (extra == "argmax": the row they name)
            lazy = w.stats if isinstance(w.stats, LazyStats) and not w.stats.resolved else None
            if extra == "argmax":
                idx = -1 if lazy is not None else int(w.stats.argmax) - row_offset
                res, self._fit_dev = ops.matnorm_update(x, nw, s32, inv_out, jitter, idx, stats=lazy)
            else:
                res, self._fit_dev = ops.matnorm_update(x, nw, s32, inv_out, jitter, -1, stats=lazy)
                res = res[:4]
            return res[:3] + (res[3] == 0,) + res[4:]
        self._fit_dev = None
        if extra is not None:
            extra = x[int(w.stats.argmax) - row_offset]
        s1 = ops.wmoments_diag(x, nw, shift1=s32)
        if not replicated:
            all_reduce_sum(s1)
        shift = (s1 if s64 is None else s1 + s64).float().reshape(H, d)
        acc = ops.wmoments_matnorm(x, nw, shift=shift, inv_cov_out=inv_out)
        if not replicated:
            all_reduce_sum(acc)
        fin = ops.finalize_matnorm(s1, acc, H, d, jitter=jitter)
        parts = [fin["mean64"].reshape(-1), fin["cov64"].reshape(-1), fin["chol64"].reshape(-1),
                 fin["status"].double()]
        if extra is not None:
            parts.append(extra.reshape(-1).double())
        packed = torch.cat(parts).cpu().numpy()
        mean = packed[:H * d].reshape(H, d)
        cov = packed[H * d:H * d + H * H].reshape(H, H)
        chol = packed[H * d + H * H:H * d + 2 * H * H].reshape(H, H)
        pd = packed[H * d + 2 * H * H] == 0.0
        if extra is not None:
            return mean, cov, chol, pd, packed[H * d + 2 * H * H + 1:].reshape(extra.shape)
        return mean, cov, chol, pd


class BaseFeatures(BasePrimitive):
    """Feature-based policies (ppi/policies.py:393-516)."""

    def update_timesteps(self, t, anneal=1.0, eps=1e-5):
        self.t = t
        if anneal < 1.0:
            self.covariance_in = anneal * self.covariance_in + (1 - anneal) * self.covariance_in_init
            self.covariance_in_sqrt = self._chol(self.covariance_in)

    def __call__(self, n_samples):
        """-> (limited trajectories ys (K,T,d or 2d), feature weights ws (K,F,d))."""
        feat_t, dfeat_t = self._device_features()
        lo, hi = _limits(self.limiter) if self.limiter else (None, None)
        z, seed, draw, begin, n_local, n_total = _draw(self.sampler, n_samples, self.dim_sample)
        ys, ws = ops.sample_features(n_local, self.mean_fn, self.mean, self.covariance_in_sqrt,
                                     self.cov_out_sqrt, feat_t, dfeat_t, z=z, seed=seed, draw_id=draw,
                                     row_offset=begin, lower=lo, upper=hi)
        return DeviceArray(ys, begin, n_total), DeviceArray(ws, begin, n_total)

    def _device_features(self):
        """Phi(t) / Phi'(t) as device tensors, rebuilt only when the time grid (its values) or the feature
        parameters change: the reference recomputes them on every call
        (ppi/policies.py:406-408), 0.45 ms of host NumPy per draw at T = 1000."""
        t = self.t
        # keyed on the whole time grid (T fp64 values): an edit of its interior rebuilds the features
        key = (t.shape, np.ascontiguousarray(t).tobytes(), getattr(self, "ls", None),
               getattr(self, "order", None), self.dim_features, bool(self.use_derivatives))
        cache = getattr(self, "_feat_cache", None)
        if cache is None or cache[0] != key:
            feat_t = f32(self.feat(t))
            dfeat_t = f32(self.dfeat(t)) if self.use_derivatives else None
            self._feat_cache = cache = (key, feat_t, dfeat_t)
        return cache[1], cache[2]

    def predict(self, only_mean=False):
        feat_t = self.feat(self.t)
        mu_y = self.limiter(self.mean_fn[None, :] + feat_t @ self.mean)
        if only_mean:
            return mu_y
        sigma_y_in = feat_t @ self.covariance_in @ feat_t.T
        sigma_y_out = self.covariance_out
        std_y_out = np.sqrt(np.outer(np.diag(sigma_y_in), np.diag(sigma_y_out)))
        return mu_y, sigma_y_in, sigma_y_out, std_y_out

    def feat(self, t):
        raise NotImplementedError

    def dfeat(self, t):
        raise NotImplementedError

    def _condition(self, t, action):
        """Bayesian linear-regression conditioning on (t, action), ppi/policies.py:449-459."""
        phi = self.feat(t)
        prior_precision = np.linalg.inv(self.covariance_in)
        precision = phi.T @ phi + prior_precision
        rhs = (action - self.mean_fn[None, :]).T @ phi + self.mean.T @ prior_precision
        self.mean = np.linalg.solve(precision, rhs.T)
        self.covariance_in = np.linalg.inv(precision)

    def weighted_update(self, log_weights, samples, update_covariance_in=True):
        """Importance-weighted ML fit of the feature weights (ppi/policies.py:461-516)."""
        w = _weights(log_weights)
        x, row_offset, n_total, replicated = _samples(samples)
        mean_, cov_in_, chol_, pd = self._mavn_update(w, x, row_offset, replicated, None, 1e-12)
        ess = float(w.stats.ess)
        if pd:
            self.mean = mean_
            if update_covariance_in:
                self.covariance_in, self.covariance_in_sqrt = cov_in_, chol_
            self._seed_fit(update_covariance_in)
        else:
            logging.warning(f"Update error, ESS={ess}")
            logging.warning("Weighted update error, reverting")
            ess = int(w.stats.n_valid)
        return ess, 0.0


class RbfFeatures(BaseFeatures):
    """Radial basis function features (ppi/policies.py:519-590)."""

    def __init__(self, time_sequence, action_dimension, mean, covariance_in, covariance_out, lengthscale,
                 n_features, sampler, use_derivatives, limiter=null_limiter, add_bias=False, *args,
                 **kwargs):
        assert lengthscale > 0.0
        assert mean.shape == (action_dimension,), f"{mean.shape} != {(action_dimension,)}"
        assert covariance_in.shape == (1,)
        self.add_bias = add_bias
        self.ls = lengthscale / np.sqrt(2)
        self.n_features = n_features
        self.dim_features = n_features + (1 if add_bias else 0)
        self.centres = np.linspace(time_sequence[0], time_sequence[-1], n_features)[:, None]
        self.norm = 1 / np.sqrt(np.sqrt(np.pi) * n_features * self.ls)
        self.mean_fn = mean
        super().__init__(time_sequence, action_dimension, np.zeros((self.dim_features, action_dimension)),
                         covariance_in.item() * np.eye(self.dim_features), covariance_out, sampler, limiter,
                         use_derivatives)

    def _bump(self, t):
        u = (t[:, None] - self.centres.T) / self.ls
        return u, self.norm * np.exp(-0.5 * u * u)

    def feat(self, t):
        cols = [self._bump(t)[1]]
        if self.add_bias:
            cols.append(np.ones((t.shape[0], 1)))
        return np.concatenate(cols, axis=1)

    def dfeat(self, t):
        u, g = self._bump(t)
        cols = [-(u / self.ls) * g]
        if self.add_bias:
            cols.append(np.zeros((t.shape[0], 1)))
        return np.concatenate(cols, axis=1)


class RffFeatures(BaseFeatures):
    """Gauss-Hermite quadrature Fourier features (ppi/policies.py:593-663)."""

    def __init__(self, time_sequence, action_dimension, mean, covariance_in, covariance_out, lengthscale,
                 order, sampler, use_derivatives, add_bias=False, limiter=null_limiter, *args, **kwargs):
        assert lengthscale > 0.0
        assert mean.shape == (action_dimension,)
        assert covariance_in.shape == (1,)
        assert covariance_out.shape == (action_dimension, action_dimension)
        self.add_bias = add_bias
        self.ls = lengthscale
        self.order = order
        nodes, weights = np.polynomial.hermite.hermgauss(2 * order)
        self.x = np.sqrt(2) * nodes[order:] / self.ls       # positive frequencies
        self.w = 2 * weights[order:] / np.sqrt(np.pi)
        self.dim_features = 2 * order + (1 if add_bias else 0)
        self.mean_fn = mean
        super().__init__(time_sequence, action_dimension, np.zeros((self.dim_features, action_dimension)),
                         covariance_in.item() * np.eye(self.dim_features), covariance_out, sampler, limiter,
                         use_derivatives)

    def feat(self, t):
        phase = np.outer(t, self.x)
        amp = np.sqrt(self.w)[None, :]
        cols = [np.cos(phase) * amp, np.sin(phase) * amp]
        if self.add_bias:
            cols.append(np.ones((t.shape[0], 1)))
        return np.concatenate(cols, axis=1)

    def dfeat(self, t):
        phase = np.outer(t, self.x)
        amp = (self.x * np.sqrt(self.w))[None, :]
        cols = [-np.sin(phase) * amp, np.cos(phase) * amp]
        if self.add_bias:
            cols.append(np.zeros((t.shape[0], 1)))
        return np.concatenate(cols, axis=1)


class BaseKernel(BasePrimitive):
    """Kernel (GP) policies (ppi/policies.py:666-860)."""

    mean_fn: np.ndarray
    covariance_in_prior: np.ndarray = None
    covariance_in_prior_inv: np.ndarray = None
    kernel_kind: str = None

    def k(self, t1, t2):
        raise NotImplementedError

    def reset_covariance(self):
        self.covariance_in = self.k(self.t, self.t)
        self.covariance_in_sqrt = self._chol(self.covariance_in)

    def __call__(self, n_samples):
        lo, hi = _limits(self.limiter)
        z, seed, draw, begin, n_local, n_total = _draw(self.sampler, n_samples, self.dim_sample)
        xs = ops.sample_matnorm(n_local, self.mean_fn, self.mean, self.covariance_in_sqrt,
                                self.cov_out_sqrt, z=z, seed=seed, draw_id=draw, row_offset=begin,
                                lower=lo, upper=hi, l_out_is_diag=self._l_out_is_diag())
        handle = DeviceArray(xs, begin, n_total)
        return handle, handle  # the clipped samples are also the fitted "parameters" (:744-745)

    def compute_prior(self, plan_time):
        self.t = plan_time
        self.covariance_in_prior = self.k(plan_time, plan_time)
        self.covariance_in_prior_inv = np.linalg.inv(self.covariance_in_prior)

    def optimize_hyperparams(self, target_matrix):
        """Kernel hyper-parameters by maximum likelihood of a target trajectory (ppi/policies.py:696-714):
        offline, SciPy L-BFGS-B on the host over the Gram matrices of the device kernel."""
        from scipy.optimize import minimize
        from scipy.stats import multivariate_normal

        target = vec(np.asarray(target_matrix, np.float64))[:, 0]
        zero = np.zeros(self.dim_features * self.dim_out)

        def objective(param):
            cov = np.kron(self.covariance_out, self._k(self.t, self.t, *param))
            return -multivariate_normal.logpdf(target, mean=zero, cov=cov)

        res = minimize(objective, x0=np.ones_like(np.array(self.param, dtype=np.float64)),
                       bounds=self.param_bounds, method="L-BFGS-B")
        logging.info(res)
        self.param = res.x

    def loglikelihood(self, x):
        """Mean matrix-normal log-likelihood of (n, H, d) samples, ppi/policies.py:716-733."""
        x = np.asarray(x, np.float64)
        n, d_i, d_o = x.shape
        assert d_i == self.dim_features and d_o == self.dim_out
        out_sqrt = np.linalg.cholesky(self.covariance_out)
        in_inv = np.linalg.solve(self.covariance_in_sqrt, self.covariance_in_sqrt.T).T
        out_inv = np.linalg.solve(out_sqrt, out_sqrt.T).T
        diff = x - self.mean[None, ...] - self.mean_fn[None, None, :]
        op = np.einsum("bij,ik,bkl->bjl", diff, in_inv, diff)
        llh = -0.5 * np.trace(np.einsum("ij,bjk->bik", out_inv, op)) / n
        llh -= 0.5 * self.dim_sample * np.log(2 * np.pi)
        llh -= 0.5 * self.dim_out * np.linalg.slogdet(self.covariance_in)[1]
        llh -= 0.5 * self.dim_features * np.linalg.slogdet(self.covariance_out)[1]
        return llh

    def timesteps_match(self, t):
        return t.shape == self.t.shape and bool((t == self.t).all())

    def _prior_device(self):
        """(prior, prior_inv) as device fp64 tensors, re-uploaded when the host arrays were replaced or edited."""
        p, pi = np.asarray(self.covariance_in_prior), np.asarray(self.covariance_in_prior_inv)
        cache = getattr(self, "_prior_dev", None)
        if (cache is None or cache[2].shape != p.shape or not np.array_equal(cache[2], p)
                or not np.array_equal(cache[3], pi)):
            self._prior_dev = cache = (f64(p), f64(pi), p.copy(), pi.copy())   # host snapshots: edits are seen
        return cache[0], cache[1]

    def update_timesteps(self, time_sequence, anneal=1.0, eps=1e-5):
        """GP shift of mean / covariance onto a new time window (ppi/policies.py:761-791)."""
        assert 0.0 <= anneal <= 1.0
        if self.timesteps_match(time_sequence):
            return
        n_new = time_sequence.shape[0]
        if n_new != self.covariance_in_prior.shape[0]:
            self.covariance_in_prior = self.k(self.t, self.t)
            self.covariance_in_prior_inv = np.linalg.inv(self.covariance_in_prior)
        if self.kernel_kind in _lib.KERNEL_KINDS and n_new <= 160:
            # one fp64 launch for the whole shift (ppi_kernel_shift): the Gram matrices, the five H x H
            # products and the factorisation never leave the device; the cached prior and its inverse
            # are uploaded only when compute_prior / a change of window length replaced them
            lo, hi = _limits(self.limiter)
            mean, cov, chol, status = ops.kernel_shift(
                self.kernel_kind, self.sigma, getattr(self, "ls", 1.0), getattr(self, "period", 1.0), self.t,
                time_sequence, *self._prior_device(), self.covariance_in, self.mean, self.mean_fn, lo, hi,
                anneal, eps * self.sigma)
            if status != 0:
                raise np.linalg.LinAlgError("Matrix is not positive definite")
            self.mean, self.covariance_in, self.covariance_in_sqrt = mean, cov, chol
            self.t = time_sequence
            self.dim_features = n_new
            return
        p_inv = self.covariance_in_prior_inv
        gain = p_inv @ (self.covariance_in_prior - self.covariance_in) @ p_inv.T
        cross = self.k(time_sequence, self.t)
        shifted = self.limiter(cross @ p_inv @ self.mean + self.mean_fn[None, :]) - self.mean_fn[None, :]
        cov_new = (self.k(time_sequence, time_sequence) - anneal * cross @ gain @ cross.T
                   + eps * self.sigma * np.eye(n_new))
        self.mean = shifted
        self.covariance_in = cov_new
        self.covariance_in_sqrt = self._chol(cov_new)
        self.t = time_sequence
        self.dim_features = n_new

    def _condition(self, t, action):
        cov_p = self.k(t, t)
        cov_tp = self.k(self.t, t)
        self.mean = cov_tp @ np.linalg.solve(cov_p, action - self.mean_fn[None, :])
        self.covariance_in = self.covariance_in - cov_tp @ np.linalg.solve(cov_p, cov_tp.T)
        self.covariance_in_sqrt = self._chol(self.covariance_in)

    def weighted_update(self, log_weights, samples, update_covariance_in=True):
        """ppi/policies.py:804-848."""
        w = _weights(log_weights)
        x, row_offset, n_total, replicated = _samples(samples)
        H, d = x.shape[1], x.shape[2]
        shift1 = np.broadcast_to(np.asarray(self.mean_fn, np.float64)[None, :], (H, d))
        if replicated or dist_info()[1] == 1:
            # the most likely sequence rides along in the one device->host read of the moments
            mean_, cov_in_, chol_, pd, best = self._mavn_update(
                w, x, row_offset, replicated, shift1, 0.0, extra="argmax")
            self.map_sequence = best
        else:
            self.map_sequence = _row(x, row_offset, n_total, replicated, int(w.stats.argmax))
            mean_, cov_in_, chol_, pd = self._mavn_update(w, x, row_offset, replicated, shift1, 0.0)
        self.mean = mean_
        ess = float(w.stats.ess)
        if update_covariance_in and pd:
            self.covariance_in, self.covariance_in_sqrt = cov_in_, chol_
        self._seed_fit(update_covariance_in and pd)
        if update_covariance_in:
            if not pd:
                logging.warning(f"Update error, ESS={ess}")
                logging.warning("Weighted update error, reverting")
                jitter = SIGMA_MIN * np.eye(self.dim_features)
                self.covariance_in = self.covariance_in + jitter
                self.covariance_in_sqrt = self.covariance_in_sqrt + jitter
                ess = int(w.stats.n_valid)
        return ess, 0.0

    def predict(self, only_mean=False):
        mu_y = self.mean_fn[None, :] + self.mean
        if only_mean:
            return mu_y
        # the reference reshapes diag(kron(cov_out, cov_in)) -- actuator-major -- row by row into
        # (H, d) (ppi/policies.py:856-859); the same element order is kept here
        sigma_y = np.sqrt(np.outer(np.diag(self.covariance_out), np.diag(self.covariance_in)).reshape(mu_y.shape))
        return mu_y, self.covariance_in, self.covariance_out, sigma_y


def _gram_host(kind, t1, t2, sigma, ls=1.0, period=1.0):
    return ops.gram(kind, t1, t2, sigma, ls, period).cpu().numpy()


class StationaryKernel(BaseKernel):
    """Squared exponential, Matern, periodic (ppi/policies.py:863-1037).  The Gram matrix and its
    Cholesky factor come from the single-block fp64 device kernel."""

    param_bounds = ((1e-5, None), (1e-3, 1e3))

    def __init__(self, time_sequence, action_dimension, mean, covariance_in, covariance_out, lengthscale,
                 sampler, limiter=null_limiter, *args, **kwargs):
        self.dim_features = time_sequence.shape[0]
        assert lengthscale > 0.0
        assert mean.shape == (action_dimension,), f"{mean.shape} != (action_dimension,)"
        assert covariance_in.shape == (1,)
        assert covariance_out.shape == (action_dimension, action_dimension)
        self.mean_fn = mean
        self.ls = lengthscale
        self.sigma = covariance_in.item()
        super().__init__(time_sequence, action_dimension, np.zeros((self.dim_features, action_dimension)),
                         self.k(time_sequence, time_sequence), covariance_out, sampler, limiter)

    @property
    def param(self):
        return self.sigma, self.ls

    @param.setter
    def param(self, params):
        self.sigma, self.ls = params

    def k(self, t1, t2):
        return self._k(t1, t2, sigma=self.sigma, ls=self.ls)


class SquaredExponentialKernel(StationaryKernel):
    kernel_kind = "se"
    can_optimize = True
    param_bounds = ((1e-5, None), (1e-5, 1e3))

    @staticmethod
    def _k(t1, t2, sigma, ls, lib=np):
        return _gram_host("se", t1, t2, sigma, ls)


class Matern12Kernel(StationaryKernel):
    kernel_kind = "matern12"
    @staticmethod
    def _k(t1, t2, sigma, ls, eps=1e-8, lib=np):
        return _gram_host("matern12", t1, t2, sigma, ls)


class Matern32Kernel(StationaryKernel):
    kernel_kind = "matern32"
    @staticmethod
    def _k(t1, t2, sigma, ls, eps=1e-8, lib=np):
        return _gram_host("matern32", t1, t2, sigma, ls)


class Matern52Kernel(StationaryKernel):
    kernel_kind = "matern52"
    @staticmethod
    def _k(t1, t2, sigma, ls, eps=1e-8, lib=np):
        return _gram_host("matern52", t1, t2, sigma, ls)


class PeriodicKernel(StationaryKernel):
    kernel_kind = "periodic"
    param_bounds = ((1e-3, None), (1e-4, 1e3), (1e-3, 1e3))

    def __init__(self, time_sequence, action_dimension, mean, covariance_in, covariance_out, lengthscale,
                 period, sampler, limiter=null_limiter, use_derivatives=False, *args, **kwargs):
        self.period = period
        super().__init__(time_sequence, action_dimension, mean, covariance_in, covariance_out, lengthscale,
                         sampler, limiter)

    @property
    def param(self):
        return self.sigma, self.ls, self.period

    @param.setter
    def param(self, params):
        self.sigma, self.ls, self.period = params

    def k(self, t1, t2):
        return self._k(t1, t2, sigma=self.sigma, ls=self.ls, per=self.period)

    @staticmethod
    def _k(t1, t2, sigma, ls, per, eps=1e-8, lib=np):
        return _gram_host("periodic", t1, t2, sigma, ls, per)


class WhiteNoiseKernel(BaseKernel):
    """ppi/policies.py:1040-1104."""

    param_bounds = ((1e-5, None),)

    def __init__(self, time_sequence, action_dimension, mean, covariance_in, covariance_out, sampler,
                 use_derivatives=False, limiter=null_limiter, *args, **kwargs):
        self.dim_features = time_sequence.shape[0]
        assert mean.shape == (action_dimension,), f"{mean.shape} != (action_dimension,)"
        assert covariance_in.shape == (1,)
        assert covariance_out.shape == (action_dimension, action_dimension)
        self.mean_fn = mean
        self.sigma = covariance_in.item()
        super().__init__(time_sequence, action_dimension, np.zeros((self.dim_features, action_dimension)),
                         self.k(time_sequence, time_sequence), covariance_out, sampler, limiter,
                         use_derivatives)

    @property
    def param(self):
        return self.sigma

    @param.setter
    def param(self, params):
        self.sigma = params[0]

    def k(self, t1, t2):
        return self._k(t1, t2, self.sigma)

    @staticmethod
    def _k(t1, t2, sigma, lib=np):
        return _gram_host("white", t1, t2, sigma)

    def update_timesteps(self, time_sequence, anneal=1.0, eps=1e-7):
        if self.timesteps_match(time_sequence):
            return
        n_new = time_sequence.shape[0]
        keep = 1.0 * (time_sequence.reshape(-1, 1) == self.t.reshape(1, -1))  # new x old selector
        fresh = np.eye(n_new) - keep @ keep.T
        self.mean = keep @ self.mean
        self.covariance_in = keep @ self.covariance_in @ keep.T + fresh @ self.k(time_sequence, time_sequence)
        self.covariance_in_sqrt = self._chol(self.covariance_in)
        self.t = time_sequence
        self.dim_features = n_new


class WhiteNoiseIid(object):
    """Independent Gaussian noise per (time step, actuator) (ppi/policies.py:1107-1195)."""

    can_optimize = False
    map_sequence: np.ndarray = None

    def __init__(self, time_sequence, action_dimension, mean, covariance_in, covariance_out,
                 use_derivatives=False, limiter=null_limiter, sampler=None, *args, **kwargs):
        self.dim_features = time_sequence.shape[0]
        self.dim_out = action_dimension
        assert mean.shape == (action_dimension,), f"{mean.shape} != (action_dimension,)"
        assert covariance_in.shape == (1,)
        assert covariance_out.shape == (action_dimension, action_dimension)
        self.t = time_sequence
        self.sigma = np.sqrt(covariance_out * covariance_in.item())
        self.mean_fn = mean
        self.mean = np.zeros((self.dim_features, action_dimension))
        self.std = np.ones((self.dim_features, action_dimension)) @ self.sigma
        self.limiter = limiter
        # the reference draws with np.random.randn directly (:1167); here a device Philox stream
        self._rng = samplers.MonteCarlo(self.dim_features * action_dimension)
        self.injected_normals = []  # parity harness: queue of (K, H, d) standard normals

    def compute_prior(self, t):
        pass

    def reset_covariance(self):
        self.std = np.ones((self.dim_features, self.dim_out)) @ self.sigma

    @property
    def covariance_out(self):
        return np.diag(self.std.mean(axis=0) ** 2)

    @property
    def entropy(self):
        # diagonal covariance: log det = sum log var, with the reference's clamp
        d = self.dim_out * self.dim_features
        logdet = float(np.sum(np.log(self.std.flatten() ** 2)))
        return 0.5 * _clamped_logdet(logdet) + (d / 2) * (1 + np.log(2 * np.pi))

    def __call__(self, n_samples):
        lo, hi = _limits(self.limiter)
        begin, end = shard_rows(n_samples)
        if self.injected_normals:
            z = f32(self.injected_normals.pop(0)).reshape(n_samples, -1)[begin:end].contiguous()
            seed = draw = 0
        else:
            z, seed, draw = self._rng.device_normals(n_samples, self.dim_features * self.dim_out)
        xs = ops.sample_iid(end - begin, self.mean_fn, self.mean, self.std, z=z, seed=seed, draw_id=draw,
                            row_offset=begin, lower=lo, upper=hi)
        handle = DeviceArray(xs, begin, n_samples)
        return handle, handle

    def weighted_update(self, log_weights, samples, update_covariance_in=True):
        w = _weights(log_weights)
        x, row_offset, n_total, replicated = _samples(samples)
        K_local, H, d = x.shape
        nw = _local(w.nw, row_offset, K_local)
        self.map_sequence = _row(x, row_offset, n_total, replicated, int(w.stats.argmax))
        base = np.ascontiguousarray(np.broadcast_to(np.asarray(self.mean_fn, np.float64)[None, :], (H, d)))
        s1 = ops.wmoments_diag(x, nw, shift1=base)
        if not replicated:
            all_reduce_sum(s1)
        if update_covariance_in:
            centre = (s1 + f64(base).reshape(-1)).float()
            _, s2 = ops.wmoments_diag(x, nw, shift2=centre, second=True)
            if not replicated:
                all_reduce_sum(s2)
            self.std = np.sqrt(s2.cpu().numpy().reshape(H, d))
        self.mean = s1.cpu().numpy().reshape(H, d)
        return float(w.stats.ess), 0.0

    def update_timesteps(self, time_sequence, anneal=1.0, eps=1e-7):
        n_new = time_sequence.shape[0]
        keep = 1.0 * (time_sequence.reshape(-1, 1) == self.t.reshape(1, -1))
        std_new = np.ones((n_new, self.dim_out)) @ self.sigma
        fresh = np.eye(n_new) - keep @ keep.T
        self.mean = keep @ self.mean
        std = np.sqrt(np.square(keep @ self.std) + np.square(fresh @ std_new))
        self.std = anneal * std + (1 - anneal) * std_new
        self.t = time_sequence
        self.dim_features = n_new

    def predict(self, only_mean=False):
        mean = self.mean_fn[None, :] + self.mean
        return mean if only_mean else (mean, self.std ** 2)


def ema_filter_matrix(n, beta):
    """The reference's `convolve(signal, [1 - beta, beta])` (ppi/policies.py:1257-1267, :1283) as a
    lower-triangular matrix: s[0] = x[0], s[t] = (1 - beta) s[t-1] + beta x[t] (the loop reads the
    already smoothed s[t-1]); identity when the horizon is not longer than the filter."""
    f = np.eye(n)
    for t in range(1, n):
        f[t] = (1.0 - beta) * f[t - 1]
        f[t, t] = beta
    return f


def powerlaw_mixing_matrix(n, beta):
    """colorednoise.powerlaw_psd_gaussian(beta, size=(.., n)) (third-party, colorednoise==2.0.0,
    ppi.yaml:23; Timmer & Koenig 1995) is a LINEAR map of its 2 (n//2 + 1) frequency-domain normals:
    y = M xi with  s_k ~ N(0, f_k^-beta) (f_0 := f_1, DC and Nyquist real and scaled by sqrt 2),
    y = irfft(s) / sigma,  sigma = 2 sqrt(sum_k w_k^2) / n.  Returns M (n, 2 m)."""
    f = np.fft.rfftfreq(n)
    f[0] = f[1]  # fmin = 1 / n
    scale = f ** (-beta / 2.0)
    w = scale[1:].copy()
    w[-1] *= (1 + (n % 2)) / 2.0
    sigma = 2.0 * np.sqrt(np.sum(w ** 2)) / n
    m = f.shape[0]
    cols = []
    for part in (0, 1):  # real draws, then imaginary draws
        for k in range(m):
            s = np.zeros(m, dtype=np.complex128)
            amp = scale[k]
            if part == 0:
                if k == 0 or (k == m - 1 and n % 2 == 0):
                    amp *= np.sqrt(2.0)
                s[k] = amp
            else:
                if k == 0 or (k == m - 1 and n % 2 == 0):
                    amp = 0.0
                s[k] = 1j * amp
            cols.append(np.fft.irfft(s, n=n) / sigma)
    return np.stack(cols, axis=1)


class _FilteredNoise(WhiteNoiseIid):
    """iid prior whose base normals pass through a fixed H x H lower-triangular map before the
    element-wise transform: both steps run on the device (ppi_sample_matnorm with L_out = I, then
    ppi_sample_iid on its output)."""

    def _mix(self):
        raise NotImplementedError

    def _base_normals(self, n_local, n_samples, begin, end):
        H, d = self.dim_features, self.dim_out
        if self.injected_normals:
            z = f32(self.injected_normals.pop(0)).reshape(n_samples, -1)[begin:end].contiguous()
            seed = draw = 0
        else:
            z, seed, draw = self._rng.device_normals(n_samples, H * d)
        return ops.sample_matnorm(n_local, np.zeros(d), np.zeros((H, d)), self._mix(), np.eye(d), z=z,
                                  seed=seed, draw_id=draw, row_offset=begin, l_out_is_diag=True)


class SmoothExplorationNoise(_FilteredNoise):
    """Exponentially smoothed exploration noise (ppi/policies.py:1270-1303): the noise, not the
    mean, goes through the causal filter [1 - beta, beta]."""

    def __init__(self, time_sequence, action_dimension, mean, covariance_in, covariance_out, sampler=None,
                 beta=2.0, limiter=null_limiter, use_derivatives=False, *args, **kwargs):
        assert 0.0 < beta < 1.0, f"beta is {beta}"
        self.beta = beta
        self.filter_coef = np.array([1 - beta, beta])
        super().__init__(time_sequence, action_dimension, mean, covariance_in, covariance_out,
                         use_derivatives, limiter)

    def _mix(self):
        return ema_filter_matrix(self.dim_features, self.beta)

    def __call__(self, n_samples):
        lo, hi = _limits(self.limiter)
        begin, end = shard_rows(n_samples)
        zs = self._base_normals(end - begin, n_samples, begin, end)
        xs = ops.sample_iid(end - begin, self.mean_fn, self.mean, self.std, z=zs.reshape(end - begin, -1),
                            lower=lo, upper=hi)
        handle = DeviceArray(xs, begin, n_samples)
        return handle, handle


class SmoothActionNoise(SmoothExplorationNoise):
    """The variant that smooths the whole action sequence, mean included (ppi/policies.py:1306-1319)."""

    def __call__(self, n_samples):
        lo, hi = _limits(self.limiter)
        begin, end = shard_rows(n_samples)
        H, d = self.dim_features, self.dim_out
        if self.injected_normals:
            z = f32(self.injected_normals.pop(0)).reshape(n_samples, -1)[begin:end].contiguous()
            seed = draw = 0
        else:
            z, seed, draw = self._rng.device_normals(n_samples, H * d)
        xs = ops.sample_iid(end - begin, self.mean_fn, self.mean, self.std, z=z, seed=seed, draw_id=draw,
                            row_offset=begin)
        smooth = ops.sample_matnorm(end - begin, np.zeros(d), np.zeros((H, d)), self._mix(), np.eye(d),
                                    z=xs.reshape(end - begin, -1), lower=lo, upper=hi, l_out_is_diag=True)
        handle = DeviceArray(smooth, begin, n_samples)
        return handle, handle


class ColouredNoise(_FilteredNoise):
    """Power-law (1/f^beta) noise over the time axis (ppi/policies.py:1198-1254).  The reference
    draws with colorednoise.powerlaw_psd_gaussian, a linear map M of frequency-domain normals; here
    the same stationary process is drawn as chol(M M^T) xi on the device (equal in distribution, not
    draw by draw: the third-party package is absent, parity is on the covariance)."""

    def __init__(self, time_sequence, action_dimension, mean, covariance_in, covariance_out, sampler=None,
                 use_derivatives=False, beta=2.0, limiter=null_limiter, *args, **kwargs):
        self.beta = beta
        self.sampler = sampler(0) if sampler is not None else samplers.MonteCarlo(0)
        super().__init__(time_sequence, action_dimension, mean, covariance_in, covariance_out,
                         use_derivatives, limiter)
        self._mix_cache = None

    def _mix(self):
        H = self.dim_features
        if self._mix_cache is None or self._mix_cache[0] != (H, self.beta):
            if H > 1:
                m = powerlaw_mixing_matrix(H, self.beta)
                c = m @ m.T
                mix = np.linalg.cholesky(0.5 * (c + c.T) + 1e-12 * np.eye(H))
            else:
                mix = np.eye(1)
            self._mix_cache = ((H, self.beta), mix)
        return self._mix_cache[1]

    def update_timesteps(self, time_sequence, anneal=1.0, eps=1e-7):
        super().update_timesteps(time_sequence, anneal)
        if isinstance(self.sampler, samplers.Particles) and self.sampler.has_particles:
            particles = np.copy(self.sampler.particles)
            last_action = particles[:, -1, :]
            particles[:, :-1, :] = self.sampler.particles[:, 1:, :]
            particles[:, -1, :] = last_action
            self.sampler.particles = particles

    def __call__(self, n_samples):
        lo, hi = _limits(self.limiter)
        begin, end = shard_rows(n_samples)
        zs = self._base_normals(end - begin, n_samples, begin, end)
        if getattr(self.sampler, "particles", None) is not None:  # ppi/policies.py:1245-1246
            parts = np.asarray(self.sampler.particles)
            n = min(parts.shape[0], n_samples)
            lo_r, hi_r = max(begin, 0), min(end, n)
            if hi_r > lo_r:
                zs[lo_r - begin:hi_r - begin] = f32(parts[lo_r:hi_r, :zs.shape[1], :])
        xs = ops.sample_iid(end - begin, self.mean_fn, self.mean, self.std, z=zs.reshape(end - begin, -1),
                            lower=lo, upper=hi)
        handle = DeviceArray(xs, begin, n_samples)
        return handle, handle


class LinearGaussianDynamicalSystemKernel(BaseKernel):
    """GP prior of a discrete-time integrator chain (ppi/policies.py:1329-1410).  The Gram matrix
    is a once-per-construction H x H host computation (block algebra on (H order)^2 matrices); the
    per-iteration sampling / update path is BaseKernel's device path."""

    can_optimize = False

    def __init__(self, time_sequence, action_dimension, mean, covariance_in, covariance_out, order, sampler,
                 use_derivatives=False, limiter=null_limiter, *args, **kwargs):
        self.dim_features = time_sequence.shape[0]
        assert order in [1, 2, 3]
        assert mean.shape == (action_dimension,)
        assert covariance_in.shape == (1,)
        self.mean_fn = mean
        self.order = order
        self.sigma = covariance_in.item()
        self.Q = np.zeros((order, order))
        self.Q[-1, -1] = self.sigma
        mean_ = np.zeros((self.dim_features, action_dimension))
        covariance_in = self.k(time_sequence, time_sequence)
        super().__init__(time_sequence, action_dimension, mean_, covariance_in, covariance_out, sampler,
                         limiter, use_derivatives)

    def k(self, t1, t2):
        from scipy.linalg import block_diag

        n, o = t1.shape[0], self.order
        a = self.transition_matrx(t1[1], t1[0], d=o)   # constant time step assumed (:1375)
        a_ = np.kron(np.eye(n), np.eye(o))
        for i in range(1, n):
            a_ += np.kron(np.eye(n, k=-i), np.linalg.matrix_power(a, i))
        q_ = block_diag(*([1e-3 * np.eye(o)] + [self.Q] * (n - 1)))
        disturbance = block_diag(*([1e-6 * np.eye(o)] * n))
        gram = a_ @ q_ @ a_.T + disturbance
        return gram[::o, ::o]

    def _condition(self, t, action):
        idx = (t == self.t).nonzero()[0]
        cov_0 = self.covariance_in
        cov_p = cov_0[np.ix_(idx, idx)]
        cov_tp = cov_0[:, idx]
        self.mean = cov_tp @ np.linalg.solve(cov_p, action - self.mean_fn[None, :])
        self.covariance_in = cov_0 - cov_tp @ np.linalg.solve(cov_p, cov_tp.T)

    @staticmethod
    def transition_matrx(t2, t1, d=3):
        a = np.eye(d)
        dt = t2 - t1
        if d == 3:
            a[0, 1] = dt
            a[0, 2] = 0.5 * dt ** 2
            a[1, 2] = dt
        elif d == 2:
            a[0, 1] = dt
        elif d != 1:
            raise ValueError("Only defined for d = 1-3")
        return a
