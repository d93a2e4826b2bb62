"""Device-memory plumbing for the host classes: torch supplies HBM allocations and CUDA streams,
the arithmetic is done by libppi_b200.so (see _lib.py).  Nothing here computes on the CPU.

K-sharding: when torch.distributed is initialised (one process per GPU), every
policy call draws K / world_size rows with a Philox row offset, costs are all-gathered so that all
ranks solve the temperature identically, and the fp64 / fp32 moment partials are all-reduced.
"""
from __future__ import annotations

import ctypes as C
import weakref

import numpy as np
import torch

from . import _lib


_cuda_ok = False


def require_cuda():
    global _cuda_ok
    if not _cuda_ok:     # checked once: torch.cuda.is_available() costs ~5 us per call on the latency-bound paths
        if not torch.cuda.is_available():
            raise _lib.PpiError("ppi_b200 needs a CUDA device (B200); there is no CPU fallback")
        _lib.load()
        _cuda_ok = True
    return torch.device("cuda", torch.cuda.current_device())


def stream_ptr():
    """Raw cudaStream_t of torch's current stream on the current device (the C getter: building a
    torch.cuda.Stream object per kernel launch cost more than the launch)."""
    return C.c_void_p(torch._C._cuda_getCurrentRawStream(torch.cuda.current_device()))


def ptr(t):
    """Device pointer of a tensor (None -> NULL)."""
    if t is None:
        return C.c_void_p(0)
    assert t.is_cuda and t.is_contiguous(), "device pointers must come from contiguous CUDA tensors"
    return C.c_void_p(t.data_ptr())


def f32(a, device=None):
    """Host array / tensor -> contiguous fp32 CUDA tensor (no copy when already one)."""
    device = device or require_cuda()
    if isinstance(a, DeviceArray):
        a = a.tensor
    if isinstance(a, torch.Tensor):
        return a.to(device=device, dtype=torch.float32).contiguous()
    return torch.as_tensor(np.ascontiguousarray(np.asarray(a, dtype=np.float32)), device=device)


_small_cache = {}


def f32_cached(a, device=None):
    """f32() for small host arrays that are re-sent on every call (policy state: mean_fn, mean, factors,
    limits): the device copy is reused while the SAME array object still holds the same numbers.  The
    whole content is compared against a host snapshot taken at upload time (<= 64 K elements, a few
    microseconds), so any in-place edit is seen -- the reference re-reads its arrays on every call."""
    if not isinstance(a, np.ndarray) or a.size == 0 or a.size > 1 << 16:
        return f32(a, device)
    device = device or require_cuda()
    key = (id(a), device.index)
    hit = _small_cache.get(key)
    if (hit is not None and hit[2]() is a and hit[0][0] == a.shape and hit[0][1] == a.dtype
            and hit[0][2] == a.tobytes()):   # the weak reference rules out a recycled id; bytes: ~1 us for 7 KB
        return hit[1]
    if len(_small_cache) > 512:
        _small_cache.clear()
    t = f32(a, device)
    try:
        ref = weakref.ref(a)
    except TypeError:      # array subclasses without weak-reference support: never cached
        return t
    _small_cache[key] = ((a.shape, a.dtype, a.tobytes()), t, ref)
    return t


def seed_f32_cache(a, t):
    """Tell f32_cached that the device tensor t already holds float32(a) -- for host mirrors that were READ from
    the device (a policy's fitted mean / factor): the next sampling call then re-uses the device copy instead of
    uploading what it just downloaded.  In-place edits of `a` are still seen (content compare)."""
    if not isinstance(a, np.ndarray) or a.size == 0 or a.size > 1 << 16 or not t.is_cuda:
        return
    try:
        ref = weakref.ref(a)
    except TypeError:
        return
    if len(_small_cache) > 512:
        _small_cache.clear()
    _small_cache[(id(a), t.device.index)] = ((a.shape, a.dtype, a.tobytes()), t, ref)


def f64(a, device=None):
    device = device or require_cuda()
    if isinstance(a, torch.Tensor):
        return a.to(device=device, dtype=torch.float64).contiguous()
    return torch.as_tensor(np.ascontiguousarray(np.asarray(a, dtype=np.float64)), device=device)


_stream_objs = {}


def sync_current_stream():
    """torch.cuda.current_stream().synchronize() without building a Stream object per call (~20 us each on the
    latency-bound paths): the objects are cached by raw handle."""
    dev = torch.cuda.current_device()
    raw = torch._C._cuda_getCurrentRawStream(dev)
    st = _stream_objs.get((dev, raw))
    if st is None:
        st = _stream_objs[(dev, raw)] = torch.cuda.current_stream()
    st.synchronize()


_pinned_staging = {}


def _pinned(nbytes):
    """One pinned staging buffer per process, grown on demand (a fresh cudaHostAlloc per read costs
    ~1 ms for a 32 MB matrix and serialises the eight processes of a node on the driver lock)."""
    buf = _pinned_staging.get("buf")
    if buf is None or buf.numel() < nbytes:
        buf = _pinned_staging["buf"] = torch.empty(max(int(nbytes), 1 << 16), dtype=torch.uint8, pin_memory=True)
    return buf


def to_host64(t):
    """Device tensor -> host fp64 NumPy array (the conversion to fp64 happens on the device; a pageable
    `.cpu()` of a 2048 x 2048 matrix costs ~15 ms).  Small results go through the process-wide pinned staging
    buffer and are copied out of it; large ones (>= 1 MB) land in a pinned tensor of their own, which torch's
    caching host allocator recycles once the previous result of that size has been dropped -- no cudaHostAlloc
    and no second host copy in a steady loop."""
    t64 = t.detach().to(torch.float64).contiguous()
    n = t64.numel()
    if 8 * n >= 1 << 20:
        host = torch.empty(t64.shape, dtype=torch.float64, pin_memory=True)
        host.copy_(t64, non_blocking=True)
        sync_current_stream()
        return host.numpy()
    stage = _pinned(8 * n)[: 8 * n].view(torch.float64)
    stage.copy_(t64.reshape(-1), non_blocking=True)
    sync_current_stream()
    return stage.numpy().reshape(tuple(t64.shape)).copy()


class DeviceArray:
    """Handle on device-resident samples / costs handed through the reference's duck-typed API.

    `policy(n)` returns these so that `weighted_update` recognises its own samples and never
    re-uploads them (the reference hands the array back untouched, ppi/algorithms.py:85-88).
    Host callables still work: np.asarray(handle) / handle[...] copy to the host on demand.
    """

    __array_priority__ = 1000

    def __init__(self, tensor, row_offset=0, global_rows=None):
        self.tensor = tensor
        self.row_offset = int(row_offset)           # first global sample index held by this rank
        self.global_rows = int(global_rows if global_rows is not None else tensor.shape[0])
        self._host = None

    @property
    def shape(self):
        return tuple(self.tensor.shape)

    @property
    def ndim(self):
        return self.tensor.dim()

    @property
    def dtype(self):
        return np.dtype(np.float64)

    def __len__(self):
        return self.tensor.shape[0]

    def numpy(self):
        if self._host is None:
            self._host = self.tensor.detach().cpu().numpy().astype(np.float64)
        return self._host

    def __array__(self, dtype=None, copy=None):
        a = self.numpy()
        return a if dtype is None else a.astype(dtype, copy=False)

    def __getitem__(self, idx):
        if isinstance(idx, (int, np.integer)):
            return self.tensor[int(idx)].detach().cpu().numpy().astype(np.float64)
        return self.numpy()[idx]

    def mean(self, *a, **k):
        return self.numpy().mean(*a, **k)

    def std(self, *a, **k):
        return self.numpy().std(*a, **k)

    def __repr__(self):
        return f"DeviceArray(shape={self.shape}, device={self.tensor.device})"


class DeferredGaussianSamples(DeviceArray):
    """Handle on a Gaussian draw that has not been launched yet.  The draw is a pure function of
    (seed, draw_id, global row) and of the policy state captured here, so WHEN it runs does not matter:

      * `f(handle)` with a fusable device cost (functions.Rastrigin / Styblinski) launches the sampler with the
        cost evaluated in its epilogue and never writes x (ppi_sample_gaussian_cost);
      * anything that touches `.tensor` (a host cost, a callback, CEM's elite rows, the x-space update)
        materialises x -- the same numbers, bit for bit, from the same Philox counters.

    Either way the base normals end up in HBM (`white`) for the whitened moment match."""

    def __init__(self, policy, version, mu_dev, chol_dev, n_local, dim, row_offset, global_rows, seed, draw_id,
                 after=None, pre=None):
        self._x = None
        self._pre = pre              # PrefetchedNormals: the store of this draw was generated ahead of time
        self._shape = (int(n_local), int(dim))
        self.row_offset, self.global_rows, self._host = int(row_offset), int(global_rows), None
        self._launch = (policy, version, mu_dev, chol_dev, int(seed), int(draw_id))
        self.white = None            # (policy, version, BaseNormals) once a kernel has produced the store
        self._after = after          # called once after the first launch (prefetch hook)

    def _done(self, base):
        if self._after is not None:
            after, self._after = self._after, None
            after(base)

    @property
    def tensor(self):
        if self._x is None:
            from . import ops

            policy, version, mu, chol, seed, draw = self._launch
            n, _ = self._shape
            if self.white is None:   # first launch: keep the base normals for the whitened update
                if self._pre is not None:
                    self._x, base = ops.sample_gaussian_pre(n, mu, chol, self._pre)
                else:
                    self._x, base = ops.sample_gaussian_keep(n, mu, chol, seed=seed, draw_id=draw,
                                                             row_offset=self.row_offset)
                self.white = (policy, version, base)
                self._done(base)
            else:                    # the fused launch skipped x: re-draw it from the same counters
                self._x = ops.sample_gaussian(n, mu, chol, seed=seed, draw_id=draw, row_offset=self.row_offset)
        return self._x

    @tensor.setter
    def tensor(self, value):
        self._x = value

    @property
    def materialised(self):
        return self._x is not None

    def fused_cost(self, fn_name):
        """Costs of a separable test function from the sampling epilogue, or None when x already exists
        (then the ordinary cost kernel reads it)."""
        if self._x is not None or self.white is not None:
            return None
        from . import ops

        policy, version, mu, chol, seed, draw = self._launch
        n, _ = self._shape
        _, base, cost = ops.sample_gaussian_cost(n, mu, chol, fn_name, seed=seed, draw_id=draw,
                                                 row_offset=self.row_offset, pre=self._pre)
        self.white = (policy, version, base)
        self._done(base)
        return cost

    @property
    def shape(self):
        return self._shape

    @property
    def ndim(self):
        return 2

    def __len__(self):
        return self._shape[0]

    def __repr__(self):
        return f"DeferredGaussianSamples(shape={self._shape}, launched={self.white is not None})"


# --------------------------------------------------------------------------------------------
# workspaces: one cached byte buffer per (tag, device), grown on demand
# --------------------------------------------------------------------------------------------
_workspaces = {}


def workspace(tag, nbytes):
    dev = require_cuda()
    key = (tag, dev.index)
    buf = _workspaces.get(key)
    if buf is None or buf.numel() < nbytes:
        _workspaces.pop(key, None)
        buf = torch.empty(max(int(nbytes), 1), dtype=torch.uint8, device=dev)
        _workspaces[key] = buf
    return buf


class LazyStats:
    """struct ppi_weight_stats still on the device: read (with a synchronisation) on first attribute
    access, or filled from bytes that came home with another result (ppi_matnorm_update)."""

    __slots__ = ("tensor", "_host", "offset")
    _SHIFTED = ("cost_min", "cost_max", "cost_mean")

    def __init__(self, tensor, offset=0.0):
        # offset: fp64 constant subtracted from host costs before the fp32 cast (algorithms._global_costs);
        # added back to the statistics that are not shift-invariant
        self.tensor, self._host, self.offset = tensor, None, float(offset)

    @property
    def resolved(self):
        return self._host is not None

    def resolve(self):
        if self._host is None:
            self._host = read_stats(self.tensor)
        return self._host

    def fill(self, raw):
        from . import _lib

        self._host = _lib.WeightStats.from_buffer_copy(raw)

    def __getattr__(self, name):
        v = getattr(self.resolve(), name)
        return v + self.offset if (name in LazyStats._SHIFTED and self.offset) else v


def read_stats(stats_tensor):
    """One D2H copy (with sync) of struct ppi_weight_stats -> _lib.WeightStats."""
    host = stats_tensor.cpu().numpy().tobytes()
    return _lib.WeightStats.from_buffer_copy(host)


# --------------------------------------------------------------------------------------------
# host-level timing scopes (bench.py): CUDA-event brackets around the steps that are not one C call --
# the collectives and the dense D x D tail.  Off unless profile_scopes(True); the C library keeps its own
# per-kernel slots (ppi_profile_enable).
# --------------------------------------------------------------------------------------------
_scopes = {"on": False, "events": {}}


def profile_scopes(on):
    _scopes["on"] = bool(on)
    _scopes["events"] = {}


class scope:
    def __init__(self, name):
        self.name = name

    def __enter__(self):
        if _scopes["on"] and torch.cuda.is_available():
            self.a = torch.cuda.Event(enable_timing=True)
            self.a.record()
        return self

    def __exit__(self, *exc):
        if _scopes["on"] and hasattr(self, "a"):
            b = torch.cuda.Event(enable_timing=True)
            b.record()
            _scopes["events"].setdefault(self.name, []).append((self.a, b))
        return False


def profile_scopes_read():
    """{name: (count, total device ms)} of the scopes recorded since profile_scopes(True)."""
    out = {}
    for name, evs in _scopes["events"].items():
        if evs:
            evs[-1][1].synchronize()
            out[name] = (len(evs), float(sum(a.elapsed_time(b) for a, b in evs)))
    return out


# --------------------------------------------------------------------------------------------
# K-sharding helpers (torch.distributed: NCCL on GPUs, gloo in the CPU tests)
# --------------------------------------------------------------------------------------------
def dist_info():
    import torch.distributed as dist

    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard_rows(n_total, rank=None, world=None):
    """Rows [begin, end) of the global sample matrix owned by `rank` (even split, remainder to the
    first ranks)."""
    if rank is None:
        rank, world = dist_info()
    base, rem = divmod(int(n_total), world)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def all_gather_rows(local, n_total):
    """Concatenate per-rank row blocks of shard_rows() order into the global (n_total, ...) tensor:
    one all_gather_into_tensor when the split is even (the configured case), a padded gather + one
    compaction otherwise."""
    import torch.distributed as dist

    rank, world = dist_info()
    if world == 1:
        return local
    local = local.contiguous()
    if n_total % world == 0:
        out = torch.empty((n_total,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
        dist.all_gather_into_tensor(out, local)
        return out
    sizes = [shard_rows(n_total, r, world) for r in range(world)]
    maxn = max(e - b for b, e in sizes)
    pad = torch.zeros((maxn,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[: local.shape[0]] = local
    out = torch.empty((world * maxn,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, pad)
    return torch.cat([out[r * maxn: r * maxn + (e - b)] for r, (b, e) in enumerate(sizes)], dim=0)


def all_reduce_sum(t):
    import torch.distributed as dist

    if dist_info()[1] > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return t


def all_reduce_max(t):
    import torch.distributed as dist

    if dist_info()[1] > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t


def broadcast_row(samples, global_index, n_total):
    """Row `global_index` of the K-sharded sample matrix, on every rank (owner broadcasts)."""
    import torch.distributed as dist

    rank, world = dist_info()
    if world == 1:
        return samples[global_index]
    owner = next(r for r in range(world) if shard_rows(n_total, r, world)[0] <= global_index
                 < shard_rows(n_total, r, world)[1])
    row = torch.empty(samples.shape[1:], dtype=samples.dtype, device=samples.device)
    if rank == owner:
        row.copy_(samples[global_index - shard_rows(n_total, owner, world)[0]])
    dist.broadcast(row, src=owner)
    return row


def gather_global_rows(samples, global_indices, row_offset, n_total):
    """Rows `global_indices` (global sample numbers, any owners) of the K-sharded sample matrix on every
    rank: each rank contributes the rows it owns, zeros elsewhere, and one all-reduce assembles them
    (the index list is short: iCEM's re-used elites, ppi/algorithms.py:162-164)."""
    idx = torch.as_tensor(global_indices, dtype=torch.int64, device=samples.device).reshape(-1)
    rank, world = dist_info()
    if world == 1:
        return samples[idx - row_offset]
    n_local = samples.shape[0]
    mine = (idx >= row_offset) & (idx < row_offset + n_local)
    out = torch.zeros((idx.numel(),) + tuple(samples.shape[1:]), dtype=samples.dtype, device=samples.device)
    if bool(mine.any()):
        out[mine] = samples[idx[mine] - row_offset]
    return all_reduce_sum(out)
