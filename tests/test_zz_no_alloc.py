"""`ppi_solve_temperature_static`: the allocation-free variant of the temperature solve (scratch in a file-scope
`static __device__` array of the library, include/ppi_b200.h) against the existing entry point
`ppi_solve_temperature` (caller-owned workspace), both called through the C ABI on the same device buffers.

Collected last on purpose (file name): a failure here must not hide other
tests.  Passes alone in a fresh process (the first call resolves the array's device address) and after the rest
of the suite (the array then holds whatever an earlier call left in it).
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
TOL = 1e-5


@pytest.fixture(scope="module")
def env():
    import torch

    assert torch.cuda.is_available()
    from ppi_b200 import _lib, device

    return torch, _lib, device


def _solve_both(env, costs, mode, hyper):
    """-> ((nw, log_w, stats) of ppi_solve_temperature, the same of ppi_solve_temperature_static)."""
    torch, _lib, device = env
    dev = device.require_cuda()
    c = torch.as_tensor(np.asarray(costs, np.float32), device=dev).contiguous()
    K = c.numel()
    out = []
    for static in (False, True):
        nw = torch.full((K,), float("nan"), dtype=torch.float32, device=dev)
        lw = torch.full((K,), float("nan"), dtype=torch.float32, device=dev)
        st = torch.zeros(_lib.STATS_BYTES, dtype=torch.uint8, device=dev)
        if static:
            _lib.call("ppi_solve_temperature_static", device.ptr(c), K, int(mode), float(hyper), device.ptr(nw),
                      device.ptr(lw), device.ptr(st), device.stream_ptr())
        else:
            nbytes = _lib.query("ppi_solve_temperature_workspace_bytes", K)
            ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
            _lib.call("ppi_solve_temperature", device.ptr(c), K, int(mode), float(hyper), device.ptr(nw),
                      device.ptr(lw), device.ptr(st), device.ptr(ws), nbytes, device.stream_ptr())
        torch.cuda.synchronize()
        out.append((nw.cpu().numpy(), lw.cpu().numpy(), device.read_stats(st)))
    return out


def _compare(ref, got, record, name):
    nw_r, lw_r, st_r = ref
    nw_g, lw_g, st_g = got
    assert np.isfinite(nw_g).all() and abs(float(nw_g.sum(dtype=np.float64)) - 1.0) < TOL
    e_nw = float(np.max(np.abs(nw_g - nw_r)) / np.max(np.abs(nw_r)))
    fin = np.isfinite(lw_r)
    assert np.array_equal(fin, np.isfinite(lw_g))                      # -inf at the same (NaN-cost) places
    e_lw = float(np.max(np.abs(lw_g[fin] - lw_r[fin])) / max(np.max(np.abs(lw_r[fin])), 1.0))
    scal = {}
    for f in ("alpha", "lse", "ess", "weighted_cost", "weight_entropy", "cost_min", "cost_max", "cost_mean",
              "cost_std", "objective"):
        a, b = getattr(st_g, f), getattr(st_r, f)
        scal[f] = abs(a - b) / max(abs(b), 1.0)
    record(name, nw=e_nw, log_w=e_lw, **scal)
    assert e_nw < TOL and e_lw < TOL, (e_nw, e_lw)
    assert max(scal.values()) < TOL, scal
    for f in ("n_valid", "argmax", "all_nan", "status"):                # index / integer work: bit-exact
        assert getattr(st_g, f) == getattr(st_r, f), f


def test_static_solver_matches_workspace_solver_on_two_calls_in_a_row(env, record):
    """Two calls in a row on the shared array, of different size, mode and launch form (one cluster of 8 CTAs,
    then the cooperative grid over 2^20 costs with NaNs): the second must not see the first one's leftovers."""
    torch, _lib, device = env
    rng = np.random.default_rng(20261105)
    c1 = rng.standard_normal(32768) * 3.0 + 10.0
    c2 = rng.gamma(2.0, 5.0, size=1 << 20)
    c2[rng.integers(0, c2.size, 1000)] = np.nan
    calls = [("ess_32k", c1, _lib.TEMP_ESS_TARGET, 512.0), ("lbps_1m_nan", c2, _lib.TEMP_LOWER_BOUND, 0.9)]
    for name, c, mode, hyper in calls:
        ref, got = _solve_both(env, c, mode, hyper)
        _compare(ref, got, record, f"solve_static[{name}]")
    # and once more the first problem: same answer as before with the second call's contents in the array
    ref, got = _solve_both(env, c1, _lib.TEMP_ESS_TARGET, 512.0)
    _compare(ref, got, record, "solve_static[ess_32k again]")


@pytest.mark.parametrize("mode,hyper", [(0, 0.9), (1, 0.01), (2, 0.5), (3, 16.0), (4, 0.5)])
def test_static_solver_every_mode_small(env, record, mode, hyper):
    """One block (K <= 4096: no barrier at all) and every temperature mode."""
    rng = np.random.default_rng(mode)
    c = rng.standard_normal(1000) ** 2
    ref, got = _solve_both(env, c, mode, hyper)
    _compare(ref, got, record, f"solve_static[mode {mode}]")


def test_static_solver_rejects_sizes_beyond_its_maximum(env):
    torch, _lib, device = env
    dev = device.require_cuda()
    small = torch.zeros(16, dtype=torch.float32, device=dev)
    st = torch.zeros(_lib.STATS_BYTES, dtype=torch.uint8, device=dev)
    k_max = 1 << 22                                                   # PPI_SOLVE_STATIC_MAX_K
    with pytest.raises(_lib.PpiError, match="exceeds the maximum"):   # refused before any pointer is touched
        _lib.call("ppi_solve_temperature_static", device.ptr(small), k_max + 1, 0, 0.9, device.ptr(small), None,
                  device.ptr(st), device.stream_ptr())
    with pytest.raises(_lib.PpiError):
        _lib.call("ppi_solve_temperature_static", None, 16, 0, 0.9, device.ptr(small), None, device.ptr(st),
                  device.stream_ptr())
    # the largest admitted size runs
    c = torch.rand(k_max, dtype=torch.float32, device=dev)
    nw = torch.empty_like(c)
    _lib.call("ppi_solve_temperature_static", device.ptr(c), k_max, 3, 4096.0, device.ptr(nw), None, device.ptr(st),
              device.stream_ptr())
    s = device.read_stats(st)
    assert abs(s.ess - 4096.0) < 1e-3 * 4096.0 and abs(float(nw.sum(dtype=torch.float64)) - 1.0) < TOL
