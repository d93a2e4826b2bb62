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


def require_cuda():
    if not torch.cuda.is_available():
        raise _lib.PpiError("ppi_b200 needs a CUDA device (B200); there is no CPU fallback")
    _lib.load()
    return torch.device("cuda", torch.cuda.current_device())


def stream_ptr():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


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
    limits): the device copy is reused while the SAME array object still holds the same numbers (shape,
    sum and end points are re-checked, so in-place edits are seen)."""
    if not isinstance(a, np.ndarray) or a.size == 0 or a.size > 1 << 16:
        return f32(a, device)
    device = device or require_cuda()
    flat = a.reshape(-1)
    sig = (a.shape, a.dtype.str, float(flat.sum()), float(flat[0]), float(flat[-1]), float(flat[a.size // 2]))
    key = (id(a), device.index)
    hit = _small_cache.get(key)
    if hit is not None and hit[0] == sig and hit[2]() is a:   # the weak reference rules out a recycled id
        return hit[1]
    if len(_small_cache) > 512:
        _small_cache.clear()
    t = f32(a, device)
    try:
        ref = weakref.ref(a)
    except TypeError:      # array subclasses without weak-reference support: never cached
        return t
    _small_cache[key] = (sig, t, ref)
    return t


def f64(a, device=None):
    device = device or require_cuda()
    if isinstance(a, torch.Tensor):
        return a.to(device=device, dtype=torch.float64).contiguous()
    return torch.as_tensor(np.ascontiguousarray(np.asarray(a, dtype=np.float64)), device=device)


def to_host64(t):
    """Device tensor -> host fp64 NumPy array through a pinned staging buffer (the conversion to
    fp64 happens on the device; a pageable `.cpu()` of a 2048 x 2048 matrix costs ~15 ms)."""
    t64 = t.detach().to(torch.float64).contiguous()
    host = torch.empty(t64.shape, dtype=torch.float64, pin_memory=True)
    host.copy_(t64, non_blocking=True)
    torch.cuda.current_stream().synchronize()
    return host.numpy()


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

    __slots__ = ("tensor", "_host")

    def __init__(self, tensor):
        self.tensor, self._host = tensor, None

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
        return getattr(self.resolve(), name)


def read_stats(stats_tensor):
    """One D2H copy (with sync) of struct ppi_weight_stats -> _lib.WeightStats."""
    host = stats_tensor.cpu().numpy().tobytes()
    return _lib.WeightStats.from_buffer_copy(host)


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
    """Concatenate per-rank row blocks of shard_rows() order into the global (n_total, ...) tensor."""
    import torch.distributed as dist

    rank, world = dist_info()
    if world == 1:
        return local
    sizes = [shard_rows(n_total, r, world) for r in range(world)]
    maxn = max(e - b for b, e in sizes)
    pad = torch.zeros((maxn,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[: local.shape[0]] = local
    out = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(out, pad)
    return torch.cat([o[: e - b] for o, (b, e) in zip(out, sizes)], dim=0)


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
