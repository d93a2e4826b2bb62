"""Tensor-level wrappers of the C ABI (include/ppi_b200.h): allocate outputs / workspaces with
torch, pass raw device pointers + the current CUDA stream, check the status.  One function per
entry point; the policy / optimizer classes and the parity tests both go through these.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from .device import f32, f32_cached, f64, ptr, read_stats, require_cuda, stream_ptr, sync_current_stream, workspace

_u64 = lambda v: C.c_uint64(int(v) & 0xFFFFFFFFFFFFFFFF)  # noqa: E731


def _opt(t, dev):
    return None if t is None else f32_cached(t, dev)


class _Keep:
    """Converts host arrays to device tensors and keeps them referenced until the C call has been
    enqueued (a temporary freed inside the argument list would hand its block to the next one)."""

    def __init__(self, dev):
        self.dev, self.alive = dev, []

    def p(self, a, cached=False):
        if a is None:
            return ptr(None)
        t = f32_cached(a, self.dev) if cached else f32(a, self.dev)
        self.alive.append(t)
        return ptr(t)


# ------------------------------------------------------------------ a3
def philox_normal(rows, cols, seed, draw_id=0, row_offset=0, full_precision=False, order="row"):
    """Base normals of a Philox draw as a (rows, cols) fp32 matrix.
    order="row": the row-quad counters of ppi_philox_normal (every sampler but the dense Gaussian one);
    order="col": the column-quad counters of the dense Gaussian sampler (D > 128), i.e. what
    sample_gaussian / _keep / _cost draw for the same (seed, draw_id, row_offset).
    full_precision (row order only): the same draw without the rounding to the fp16 grid."""
    dev = require_cuda()
    out = torch.empty((rows, cols), dtype=torch.float32, device=dev)
    assert order in ("row", "col") and not (full_precision and order == "col")
    name = "ppi_philox_normal_cq" if order == "col" else ("ppi_philox_normal_full" if full_precision
                                                          else "ppi_philox_normal")
    _lib.call(name, ptr(out), rows, cols, _u64(seed), _u64(draw_id), _u64(row_offset), stream_ptr())
    return out


# ------------------------------------------------------------------ a22
def gram(kind, t1, t2, sigma, ls=1.0, period=1.0, chol=False):
    """Kernel Gram matrix in fp64 (and its Cholesky factor + status when chol=True)."""
    dev = require_cuda()
    t1d, t2d = f64(np.asarray(t1).reshape(-1), dev), f64(np.asarray(t2).reshape(-1), dev)
    n1, n2 = t1d.numel(), t2d.numel()
    g = torch.empty((n1, n2), dtype=torch.float64, device=dev)
    if not chol:
        _lib.call("ppi_kernel_gram_chol", ptr(t1d), n1, ptr(t2d), n2, _lib.KERNEL_KINDS[kind],
                  float(sigma), float(ls), float(period), ptr(g), None, None, None, None, stream_ptr())
        return g
    l = torch.empty((n1, n1), dtype=torch.float64, device=dev)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    _lib.call("ppi_kernel_gram_chol", ptr(t1d), n1, ptr(t2d), n2, _lib.KERNEL_KINDS[kind],
              float(sigma), float(ls), float(period), ptr(g), None, ptr(l), None, ptr(status),
              stream_ptr())
    return g, l, int(status.item())


def sobol_normal(rows, cols, sv, shift, bits, first_index, shrinkage=0.9999):
    """Scrambled-Sobol base normals on the device (ppi/samplers.py:31-37); sv (cols, bits) / shift (cols)
    are uint32 device tensors holding the state of a scipy.stats.qmc.Sobol engine."""
    dev = require_cuda()
    out = torch.empty((rows, cols), dtype=torch.float32, device=dev)
    _lib.call("ppi_sobol_normal", ptr(out), rows, cols, ptr(sv), ptr(shift), int(bits), _u64(first_index),
              float(shrinkage), stream_ptr())
    return out


def cubature_points(d):
    """sqrt(d) [I; -I] as a (2 d, d) device tensor (ppi/samplers.py:47-53)."""
    dev = require_cuda()
    out = torch.empty((2 * d, d), dtype=torch.float32, device=dev)
    _lib.call("ppi_cubature_points", ptr(out), int(d), stream_ptr())
    return out


_shift_staging = {}


def kernel_shift(kind, sigma, ls, period, t_old, t_new, prior, prior_inv, cov_in, mean, mean_fn, lower,
                 upper, anneal, jitter):
    """MPC time shift of a kernel prior (BaseKernel.update_timesteps, ppi/policies.py:761-791) in one
    launch: one packed host->device copy of the window's moments, one device->host read of the result,
    both through pinned staging buffers kept per window shape.
    prior / prior_inv are device fp64 tensors (cached by the policy between control steps).
    -> (mean_new, cov_new, chol_new, status) as host fp64 arrays."""
    dev = require_cuda()
    n_old, n_new, d = t_old.shape[0], t_new.shape[0], mean.shape[1]
    has_lim = lower is not None
    sizes = [n_old, n_new, n_old * n_old, n_old * d, d] + ([d, d] if has_lim else [])
    key = (dev, n_old, n_new, d, has_lim)
    st = _shift_staging.get(key)
    if st is None:
        offs = np.cumsum([0] + sizes)
        n_out = n_new * d + 2 * n_new * n_new + 1
        h_in = torch.empty(int(offs[-1]), dtype=torch.float64, pin_memory=True)
        h_out = torch.empty(n_out, dtype=torch.float64, pin_memory=True)
        d_in = torch.empty(int(offs[-1]), dtype=torch.float64, device=dev)
        d_out = torch.empty(n_out, dtype=torch.float64, device=dev)
        ws = torch.empty(max(1, _lib.query("ppi_kernel_shift_workspace_bytes", n_old, n_new) // 8),
                         dtype=torch.float64, device=dev)
        a, b = n_new * d, n_new * d + n_new * n_new
        st = _shift_staging[key] = dict(
            h_in=h_in, h_out=h_out, d_in=d_in, d_out=d_out, ws=ws, np_in=h_in.numpy(), np_out=h_out.numpy(),
            offs=[int(o) for o in offs], in_ptrs=[d_in.data_ptr() + 8 * int(o) for o in offs[:-1]],
            out_ptrs=[d_out.data_ptr(), d_out.data_ptr() + 8 * a, d_out.data_ptr() + 8 * b,
                      d_out.data_ptr() + 8 * (n_out - 1)], cut=(a, b, b + n_new * n_new))
    buf, offs = st["np_in"], st["offs"]
    parts = [t_old, t_new, cov_in, mean, mean_fn] + ([lower, upper] if has_lim else [])
    for i, part in enumerate(parts):
        buf[offs[i]:offs[i + 1]] = np.asarray(part, np.float64).reshape(-1)
    st["d_in"].copy_(st["h_in"], non_blocking=True)
    ip, op = st["in_ptrs"], st["out_ptrs"]
    _lib.call("ppi_kernel_shift", _lib.KERNEL_KINDS[kind], float(sigma), float(ls), float(period),
              ip[0], n_old, ip[1], n_new, ptr(prior), ptr(prior_inv), ip[2], ip[3], d, ip[4],
              ip[5] if has_lim else None, ip[6] if has_lim else None, float(anneal), float(jitter),
              op[0], op[1], op[2], op[3], ptr(st["ws"]), stream_ptr())
    st["h_out"].copy_(st["d_out"], non_blocking=True)
    sync_current_stream()
    host = st["np_out"]
    a, b, c = st["cut"]
    return (host[:a].reshape(n_new, d).copy(), host[a:b].reshape(n_new, n_new).copy(),
            host[b:c].reshape(n_new, n_new).copy(), int(host[-1]))


def cholesky_small(a):
    """np.linalg.cholesky of a small fp64 matrix on the device; returns (L fp64, status)."""
    dev = require_cuda()
    a = f64(a, dev)
    n = a.shape[0]
    l = torch.empty_like(a)
    status = torch.zeros(1, dtype=torch.int32, device=dev)
    _lib.call("ppi_cholesky_small", ptr(a), n, ptr(l), None, ptr(status), stream_ptr())
    return l, int(status.item())


# ------------------------------------------------------------------ a4-a8
def _is_diag(m):
    m = np.asarray(m)
    return bool(np.count_nonzero(m - np.diag(np.diag(m))) == 0)


def sample_matnorm(K, mean_fn, mean, l_in, l_out, z=None, seed=0, draw_id=0, row_offset=0,
                   lower=None, upper=None, l_out_is_diag=False):
    dev = require_cuda()
    kp = _Keep(dev)
    mean = f32_cached(mean, dev)
    H, d = mean.shape
    x = torch.empty((K, H, d), dtype=torch.float32, device=dev)
    zt = None if z is None else f32(z, dev).reshape(K, H * d)
    lo, hi = _opt(lower, dev), _opt(upper, dev)
    _lib.call("ppi_sample_matnorm", ptr(zt), _u64(seed), _u64(draw_id), _u64(row_offset), K, H, d,
              kp.p(mean_fn, cached=True), ptr(mean), kp.p(l_in, cached=True), kp.p(l_out, cached=True),
              int(bool(l_out_is_diag)), ptr(lo), ptr(hi), ptr(x), stream_ptr())
    return x


def sample_features(K, mean_fn, mean, l_in, l_out, feat, dfeat=None, z=None, seed=0, draw_id=0,
                    row_offset=0, lower=None, upper=None):
    dev = require_cuda()
    kp = _Keep(dev)
    mean = f32_cached(mean, dev)
    F, d = mean.shape
    feat = f32(feat, dev)
    T = feat.shape[0]
    dy = 2 * d if dfeat is not None else d
    w = torch.empty((K, F, d), dtype=torch.float32, device=dev)
    y = torch.empty((K, T, dy), dtype=torch.float32, device=dev)
    zt = None if z is None else f32(z, dev).reshape(K, F * d)
    lo, hi = _opt(lower, dev), _opt(upper, dev)
    _lib.call("ppi_sample_features", ptr(zt), _u64(seed), _u64(draw_id), _u64(row_offset), K, F, d, T,
              kp.p(mean_fn, cached=True), ptr(mean), kp.p(l_in, cached=True), kp.p(l_out, cached=True), ptr(feat),
              kp.p(dfeat), ptr(lo), ptr(hi), ptr(w), ptr(y), stream_ptr())
    return y, w


def sample_iid(K, mean_fn, mean, std, z=None, seed=0, draw_id=0, row_offset=0, lower=None,
               upper=None):
    dev = require_cuda()
    kp = _Keep(dev)
    mean = f32_cached(mean, dev)
    H, d = mean.shape
    x = torch.empty((K, H, d), dtype=torch.float32, device=dev)
    zt = None if z is None else f32(z, dev).reshape(K, H * d)
    lo, hi = _opt(lower, dev), _opt(upper, dev)
    _lib.call("ppi_sample_iid", ptr(zt), _u64(seed), _u64(draw_id), _u64(row_offset), K, H, d,
              kp.p(mean_fn, cached=True), ptr(mean), kp.p(std, cached=True), ptr(lo), ptr(hi), ptr(x),
              stream_ptr())
    return x


def sample_gaussian(K, mu, L, z=None, seed=0, draw_id=0, row_offset=0, out=None):
    """x = mu + z L^T with L lower triangular (Cholesky form, ppi/samplers.py:38-39)."""
    dev = require_cuda()
    kp = _Keep(dev)
    mu = f32(mu, dev)
    D = mu.numel()
    x = out if out is not None else torch.empty((K, D), dtype=torch.float32, device=dev)
    zt = None if z is None else f32(z, dev).reshape(K, D)
    nbytes = _lib.query("ppi_sample_gaussian_workspace_bytes", K, D)
    ws = workspace("sample_gaussian", nbytes) if nbytes else None
    _lib.call("ppi_sample_gaussian", ptr(zt), _u64(seed), _u64(draw_id), _u64(row_offset), K, D,
              ptr(mu), kp.p(L), ptr(x), ptr(ws), nbytes, stream_ptr())
    return x


class BaseNormals:
    """Base-normal store of one sample_gaussian(keep=True) draw (ppi_sample_gaussian_keep): the
    tensor-core operand form of z, kept in HBM for the whitened moment match."""

    def __init__(self, store, rows, dim, injected):
        self.store, self.rows, self.dim, self.injected = store, rows, dim, bool(injected)
        self.valid = True   # cleared when a later prefetch recycles the store


def sample_gaussian_keep(K, mu, L, z=None, seed=0, draw_id=0, row_offset=0):
    """sample_gaussian + the base-normal store: -> (x, BaseNormals).  D > 128, FP16X2 scheme."""
    dev = require_cuda()
    kp = _Keep(dev)
    mu = f32(mu, dev)
    D = mu.numel()
    x = torch.empty((K, D), dtype=torch.float32, device=dev)
    zt = None if z is None else f32(z, dev).reshape(K, D)
    nbytes = _lib.query("ppi_sample_gaussian_workspace_bytes", K, D)
    ws = workspace("sample_gaussian", nbytes)
    sbytes = _lib.query("ppi_zstore_bytes", K, D, int(zt is not None))
    raw = torch.empty(sbytes + 1024, dtype=torch.uint8, device=dev)
    off = (-raw.data_ptr()) % 1024  # the store header / TMA rows want 1 KB alignment
    store = raw[off:off + sbytes]
    _lib.call("ppi_sample_gaussian_keep", ptr(zt), _u64(seed), _u64(draw_id), _u64(row_offset), K, D,
              ptr(mu), kp.p(L), ptr(x), ptr(store), sbytes, ptr(ws), nbytes, stream_ptr())
    return x, BaseNormals(store, K, D, zt is not None)


def sample_gaussian_cost(K, mu, L, fn_name, z=None, seed=0, draw_id=0, row_offset=0, want_x=False, pre=None):
    """sample_gaussian_keep with a separable test function evaluated in the sampling epilogue
    (ppi_sample_gaussian_cost): -> (x or None, BaseNormals, cost (K,) fp32).  want_x=False never writes x.
    pre: a PrefetchedNormals whose store already holds this draw (only the GEMMs run)."""
    dev = require_cuda()
    kp = _Keep(dev)
    mu = f32(mu, dev)
    D = mu.numel()
    x = torch.empty((K, D), dtype=torch.float32, device=dev) if want_x else None
    zt = None if z is None else f32(z, dev).reshape(K, D)
    nbytes = _lib.query("ppi_sample_gaussian_cost_workspace_bytes", K, D)
    ws = workspace("sample_gaussian", nbytes)
    if pre is not None:
        assert zt is None
        torch.cuda.current_stream().wait_event(pre.done)
        store, sbytes = pre.store, pre.store.numel()
    else:
        sbytes = _lib.query("ppi_zstore_bytes", K, D, int(zt is not None))
        store = _aligned_bytes(sbytes, dev)
    cost = torch.empty((K,), dtype=torch.float32, device=dev)
    _lib.call("ppi_sample_gaussian_cost", ptr(zt), _u64(seed), _u64(draw_id), _u64(row_offset), K, D, ptr(mu),
              kp.p(L), ptr(x), ptr(store), sbytes, int(pre is not None), _lib.FN_IDS[fn_name], ptr(cost), ptr(ws),
              nbytes, stream_ptr())
    return x, BaseNormals(store, K, D, zt is not None), cost


def _aligned_bytes(nbytes, dev):
    raw = torch.empty(nbytes + 1024, dtype=torch.uint8, device=dev)
    off = (-raw.data_ptr()) % 1024
    return raw[off:off + nbytes]


class PrefetchedNormals:
    """Base normals of one Philox draw generated ahead of the sampling call (ppi_philox_prefetch) on a
    side stream: the K-major store and the event that marks it complete.  `key` identifies the draw."""

    def __init__(self, key, store, done):
        self.key, self.store, self.done = key, store, done


def philox_prefetch(K, D, seed, draw_id, row_offset, store, side, key):
    """Enqueue the generation of draw (seed, draw_id) rows [row_offset, row_offset + K) on `side` after
    everything already enqueued on the current stream.  store: persistent 1 KB-aligned buffer."""
    main = torch.cuda.current_stream()
    fork = torch.cuda.Event()
    fork.record(main)
    side.wait_event(fork)
    _lib.call("ppi_philox_prefetch", _u64(seed), _u64(draw_id), _u64(row_offset), K, D, ptr(store), store.numel(),
              C.c_void_p(side.cuda_stream))
    done = torch.cuda.Event()
    done.record(side)
    return PrefetchedNormals(key, store, done)


def sample_gaussian_pre(K, mu, L, pre):
    """sample_gaussian_keep for a prefetched draw: only the GEMM passes run.  -> (x, BaseNormals)."""
    dev = require_cuda()
    kp = _Keep(dev)
    mu = f32(mu, dev)
    D = mu.numel()
    x = torch.empty((K, D), dtype=torch.float32, device=dev)
    nbytes = _lib.query("ppi_sample_gaussian_workspace_bytes", K, D)
    ws = workspace("sample_gaussian", nbytes)
    torch.cuda.current_stream().wait_event(pre.done)
    _lib.call("ppi_sample_gaussian_pre", K, D, ptr(mu), kp.p(L), ptr(x), ptr(pre.store), ptr(ws), nbytes,
              stream_ptr())
    return x, BaseNormals(pre.store, K, D, False)


def wmoments_white(base, nw, s1=None, cov=None, accumulate=False, s2=None):
    """Whitened moments of a BaseNormals store: s1 (D fp64) += sum nw z, cov (D, D) = sum nw z z^T.
    s2 (D fp64, optional) += diag of that sum with round-to-nearest products (the accurate trace)."""
    dev = require_cuda()
    kp = _Keep(dev)
    K, D = base.rows, base.dim
    s1 = torch.zeros(D, dtype=torch.float64, device=dev) if s1 is None else s1
    cov = torch.empty((D, D), dtype=torch.float32, device=dev) if cov is None else cov
    nbytes = _lib.query("ppi_wmoments_white_workspace_bytes", K, D, int(base.injected))
    ws = workspace("wmoments_dense", nbytes)
    _lib.call("ppi_wmoments_white", ptr(base.store), int(base.injected), kp.p(nw), K, D, ptr(s1), ptr(s2), ptr(cov),
              int(bool(accumulate)), ptr(ws), nbytes, stream_ptr())
    return s1, cov


# ------------------------------------------------------------------ a9
def cost_function(name, x, A=None, noise=None):
    dev = require_cuda()
    kp = _Keep(dev)
    x = f32(x, dev)
    K, D = x.shape
    c = torch.empty((K,), dtype=torch.float32, device=dev)
    _lib.call("ppi_cost_function", _lib.FN_IDS[name], ptr(x), K, D, kp.p(A),
              kp.p(noise), ptr(c), stream_ptr())
    return c


def cost_tracking(x, target):
    dev = require_cuda()
    x = f32(x, dev)
    K, T, dy = x.shape
    target = f32(target, dev)
    dt = target.shape[1]
    c = torch.empty((K,), dtype=torch.float32, device=dev)
    _lib.call("ppi_cost_tracking", ptr(x), K, T, dy, ptr(target), dt, ptr(c), stream_ptr())
    return c


# ------------------------------------------------------------------ a2, a10-a15
def _stats_tensor(dev):
    return torch.zeros(_lib.STATS_BYTES, dtype=torch.uint8, device=dev)


def solve_temperature(costs, mode, hyper, want_log_w=True, sync=True):
    """-> (nw, log_w, stats).  stats is a _lib.WeightStats (host) when sync else the device bytes."""
    dev = require_cuda()
    c = f32(costs, dev).reshape(-1)
    K = c.numel()
    nw = torch.empty_like(c)
    lw = torch.empty_like(c) if want_log_w else None
    st = _stats_tensor(dev)
    nbytes = _lib.query("ppi_solve_temperature_workspace_bytes", K)
    ws = workspace("solve", nbytes)
    _lib.call("ppi_solve_temperature", ptr(c), K, int(mode), float(hyper), ptr(nw), ptr(lw), ptr(st),
              ptr(ws), nbytes, stream_ptr())
    return nw, lw, (read_stats(st) if sync else st)


def normalise_log_weights(log_w, sync=True):
    dev = require_cuda()
    lw = f32(log_w, dev).reshape(-1)
    K = lw.numel()
    nw = torch.empty_like(lw)
    st = _stats_tensor(dev)
    nbytes = _lib.query("ppi_solve_temperature_workspace_bytes", K)
    ws = workspace("solve", nbytes)
    _lib.call("ppi_normalise_log_weights", ptr(lw), K, ptr(nw), ptr(st), ptr(ws), nbytes, stream_ptr())
    return nw, (read_stats(st) if sync else st)


def cem_weights(costs, n_elites, sync=True):
    dev = require_cuda()
    c = f32(costs, dev).reshape(-1)
    K = c.numel()
    nw = torch.empty_like(c)
    idx = torch.empty((K,), dtype=torch.int64, device=dev)
    st = _stats_tensor(dev)
    nbytes = _lib.query("ppi_cem_workspace_bytes", K)
    ws = workspace("cem", nbytes)
    _lib.call("ppi_cem_weights", ptr(c), K, int(n_elites), ptr(idx), ptr(nw), ptr(st), ptr(ws), nbytes,
              stream_ptr())
    return idx, nw, (read_stats(st) if sync else st)


# ------------------------------------------------------------------ a16-a21
def wmoments_diag(x, nw, shift1=None, shift2=None, second=False, s1=None, s2=None, absdev=None):
    """fp64 accumulators: s1 = sum nw (x - shift1), s2 = sum nw (x - shift2)^2 over rows of x (K, D).
    absdev: optional zeroed fp32 device scalar that receives max |x - shift1| over weighted rows."""
    dev = require_cuda()
    kp = _Keep(dev)
    x = f32(x, dev)
    K = x.shape[0]
    x2 = x.reshape(K, -1)
    D = x2.shape[1]
    s1 = torch.zeros(D, dtype=torch.float64, device=dev) if s1 is None else s1
    if second and s2 is None:
        s2 = torch.zeros(D, dtype=torch.float64, device=dev)
    _lib.call("ppi_wmoments_diag", ptr(x2), kp.p(nw), K, D, kp.p(shift1),
              kp.p(shift2), ptr(s1), ptr(s2 if second else None), ptr(absdev), stream_ptr())
    return (s1, s2) if second else s1


def wmoments_matnorm(x, nw, shift=None, inv_cov_out=None, acc=None):
    dev = require_cuda()
    kp = _Keep(dev)
    x = f32(x, dev)
    K, H, d = x.shape
    acc = torch.zeros((H, H), dtype=torch.float64, device=dev) if acc is None else acc
    _lib.call("ppi_wmoments_matnorm", ptr(x), kp.p(nw), K, H, d, kp.p(shift),
              kp.p(inv_cov_out), ptr(acc), stream_ptr())
    return acc


def wmoments_dense(x, nw, shift=None, cov=None, accumulate=False, dev_bound=None):
    dev = require_cuda()
    kp = _Keep(dev)
    x = f32(x, dev)
    K, D = x.shape
    cov = torch.empty((D, D), dtype=torch.float32, device=dev) if cov is None else cov
    nbytes = _lib.query("ppi_wmoments_dense_workspace_bytes", K, D)
    ws = workspace("wmoments_dense", nbytes)
    _lib.call("ppi_wmoments_dense", ptr(x), kp.p(nw), K, D, kp.p(shift), ptr(dev_bound), ptr(cov),
              int(bool(accumulate)), ptr(ws), nbytes, stream_ptr())
    return cov


def finalize_gaussian(s1=None, cov=None, shift=None, diagonal_only=False, want_mu=True,
                      want_sigma=True):
    dev = require_cuda()
    kp = _Keep(dev)
    D = (s1.numel() if s1 is not None else cov.shape[0])
    mu = torch.empty(D, dtype=torch.float32, device=dev) if want_mu else None
    sigma = torch.empty((D, D), dtype=torch.float32, device=dev) if want_sigma else None
    _lib.call("ppi_finalize_gaussian", ptr(s1), ptr(cov), D, kp.p(shift),
              int(bool(diagonal_only)), ptr(mu), ptr(sigma), stream_ptr())
    return mu, sigma


_matnorm_ws_bytes = {}


def matnorm_update(x, nw, shift1, inv_cov_out, jitter=0.0, map_index=-1, stats=None):
    """One-rank matrix-normal moment match in a single call (ppi_matnorm_update): -> (host, dev) with
    host = fp64 (mean, cov_in, chol, status[, x[map_index]]) after ONE device->host read through pinned staging,
    dev = (mean32, chol32) fp32 device copies of the fit for the next sampling call.
    stats: a LazyStats whose struct is still on the device -- the sample row is then x[stats.argmax] and
    the statistics are filled from the same read (no synchronisation since the temperature solve)."""
    from .device import _pinned

    dev = require_cuda()
    K, H, d = x.shape
    on_dev = stats is not None and not stats.resolved
    want_row = map_index >= 0 or on_dev
    n_stats = _lib.STATS_BYTES // 8 if on_dev else 0
    n = H * d + 2 * H * H + 1 + (H * d if want_row else 0) + n_stats
    packed = torch.empty(n, dtype=torch.float64, device=dev)
    mean32 = torch.empty((H, d), dtype=torch.float32, device=dev)
    chol32 = torch.empty((H, H), dtype=torch.float32, device=dev)
    nbytes = _matnorm_ws_bytes.get((H, d))
    if nbytes is None:
        nbytes = _matnorm_ws_bytes[(H, d)] = _lib.query("ppi_matnorm_update_workspace_bytes", H, d)
    ws = workspace("matnorm_update", nbytes)
    _lib.call("ppi_matnorm_update", ptr(x), ptr(nw), K, H, d, ptr(shift1), ptr(inv_cov_out), float(jitter),
              int(map_index), ptr(stats.tensor) if on_dev else None, ptr(packed), ptr(mean32), ptr(chol32), ptr(ws),
              nbytes, stream_ptr())
    stage = _pinned(8 * n)[: 8 * n].view(torch.float64)
    stage.copy_(packed, non_blocking=True)
    sync_current_stream()
    host = stage.numpy().copy()
    if on_dev:
        stats.fill(host[n - n_stats:].tobytes())
    a, b, c = H * d, H * d + H * H, H * d + 2 * H * H
    out = (host[:a].reshape(H, d), host[a:b].reshape(H, H), host[b:c].reshape(H, H), int(host[c]))
    return out + ((host[c + 1:c + 1 + H * d].reshape(H, d),) if want_row else ()), (mean32, chol32)


def finalize_matnorm(s1, acc, H, d, jitter=0.0):
    """-> dict(mean64, mean32, cov64, cov32, chol64, chol32, status) (status is a device int32)."""
    dev = require_cuda()
    out = dict(
        mean64=torch.empty((H, d), dtype=torch.float64, device=dev),
        mean32=torch.empty((H, d), dtype=torch.float32, device=dev),
    )
    if acc is not None:
        out.update(
            cov64=torch.empty((H, H), dtype=torch.float64, device=dev),
            cov32=torch.empty((H, H), dtype=torch.float32, device=dev),
            chol64=torch.empty((H, H), dtype=torch.float64, device=dev),
            chol32=torch.empty((H, H), dtype=torch.float32, device=dev),
            status=torch.zeros(1, dtype=torch.int32, device=dev),
        )
    _lib.call("ppi_finalize_matnorm", ptr(s1), ptr(acc), H, d, float(jitter), ptr(out["mean64"]),
              ptr(out["mean32"]), ptr(out.get("cov64")), ptr(out.get("cov32")), ptr(out.get("chol64")),
              ptr(out.get("chol32")), ptr(out.get("status")), stream_ptr())
    return out


# ------------------------------------------------------------------ f3: dense D x D fp64 tail
def cholesky_blocked(a, want32=False):
    """np.linalg.cholesky of a (D, D) fp64 device matrix, any D (blocked, fp64 tensor cores).
    -> (L fp64, L fp32 or None, scalars): scalars = device tensor [status, log det, -, -]."""
    dev = require_cuda()
    a = f64(a, dev)
    D = a.shape[0]
    l64 = torch.empty_like(a)
    l32 = torch.empty((D, D), dtype=torch.float32, device=dev) if want32 else None
    scal = torch.empty(4, dtype=torch.float64, device=dev)
    nbytes = _lib.query("ppi_dense_workspace_bytes", D)
    ws = workspace("dense64", nbytes)
    _lib.call("ppi_cholesky_blocked", ptr(a), D, ptr(l64), ptr(l32), ptr(scal), ptr(ws), nbytes, stream_ptr())
    return l64, l32, scal


def white_finish(s, m, l_old64, mu_old32, diag=None):
    """Tail of the whitened update in one call (ppi_white_finish).
    -> (l_new64, l_new32, mu_new32, scalars [status, log det C, tr C, |m|^2] on the device)."""
    dev = require_cuda()
    D = m.numel()
    l_new64 = torch.zeros((D, D), dtype=torch.float64, device=dev)
    l_new32 = torch.zeros((D, D), dtype=torch.float32, device=dev)
    mu_new = torch.empty(D, dtype=torch.float32, device=dev)
    scal = torch.empty(4, dtype=torch.float64, device=dev)
    nbytes = _lib.query("ppi_dense_workspace_bytes", D)
    ws = workspace("dense64", nbytes)
    _lib.call("ppi_white_finish", ptr(s), ptr(m), ptr(diag), ptr(l_old64), ptr(mu_old32), D, ptr(l_new64), ptr(l_new32),
              ptr(mu_new), ptr(scal), ptr(ws), nbytes, stream_ptr())
    return l_new64, l_new32, mu_new, scal


def tri_gram(l64, want32=False):
    """sigma = L L^T from a lower-triangular fp64 factor (ppi_tri_gram)."""
    dev = require_cuda()
    D = l64.shape[0]
    out64 = torch.empty((D, D), dtype=torch.float64, device=dev)
    out32 = torch.empty((D, D), dtype=torch.float32, device=dev) if want32 else None
    _lib.call("ppi_tri_gram", ptr(l64), D, ptr(out64), ptr(out32), stream_ptr())
    return (out64, out32) if want32 else out64


def tri_solve_sumsq(l64, b_lower64=None, v64=None):
    """(|L^-1 B|_F^2, |L^-1 v|^2) as a 2-element fp64 device tensor (ppi_tri_solve_sumsq): the solve-dependent
    terms of the Gaussian KL from Cholesky factors, without forming any inverse."""
    dev = require_cuda()
    D = l64.shape[0]
    out = torch.empty(2, dtype=torch.float64, device=dev)
    nbytes = _lib.query("ppi_dense_workspace_bytes", D)
    ws = workspace("dense64", nbytes)
    _lib.call("ppi_tri_solve_sumsq", ptr(l64), D, ptr(b_lower64), ptr(v64), ptr(out), ptr(ws), nbytes, stream_ptr())
    return out


def tri_matvec(l64, v64, x0_32):
    """x0 + L v (fp32 result) for a lower-triangular fp64 L: the mean step of the whitened update alone."""
    dev = require_cuda()
    D = v64.numel()
    out = torch.empty(D, dtype=torch.float32, device=dev)
    _lib.call("ppi_tri_matvec", ptr(l64), D, ptr(v64), ptr(x0_32), ptr(out), stream_ptr())
    return out
