"""Distribution tests of the on-device base normals (ppi_philox_normal: Philox4x32-10 + Box-Muller, values
rounded to the IEEE-half grid so that each is ONE exact tensor-core operand; ppi_philox_normal_full: the same
draw unrounded).  They replace `np.random.randn` (ppi/samplers.py:18-20), so what is pinned is the LAW:

  * chi-square of 2^26 draws against Phi on 81 cells of |z| -- the cell edges sit on MIDPOINTS between adjacent
    half-precision values, so the cell probabilities of the rounded variable are exact Phi differences;
  * tail counts beyond 4 and 5 sigma, sign symmetry, first four moments;
  * the half-precision subnormal region around zero (|z| < 2^-14) is populated at the right rate and nothing
    is flushed to zero wholesale;
  * the rounded draw is the nearest half-precision value of the unrounded one (|difference| <= ulp / 2).

All statistics are computed with torch on the device (test infrastructure).
"""
import math

import numpy as np
import pytest
from scipy.stats import norm

pytestmark = pytest.mark.gpu
ROWS, COLS = 1 << 14, 1 << 12          # 2^26 normals


@pytest.fixture(scope="module")
def draws():
    import torch

    assert torch.cuda.is_available()
    from ppi_b200 import ops

    z = ops.philox_normal(ROWS, COLS, seed=20261105, draw_id=7)
    zf = ops.philox_normal(ROWS, COLS, seed=20261105, draw_id=7, full_precision=True)
    return torch, z.reshape(-1), zf.reshape(-1)


def _half_ulp(g):
    """Half the spacing of the half-precision grid at magnitude g (g > 0 a grid point below 8)."""
    e = math.floor(math.log2(g)) if g >= 2.0 ** -14 else -14
    return 2.0 ** (e - 10) / 2.0


def _chi_square(torch, v_abs, edges):
    """Cells (edges[i], edges[i+1]) of |z| (last cell open-ended): -> (statistic, counts, expected)."""
    n = v_abs.numel()
    e = torch.as_tensor(edges, dtype=torch.float64, device=v_abs.device)
    idx = torch.bucketize(v_abs.double(), e, right=True) - 1          # edges[0] = 0 -> every value lands >= 0
    counts = torch.bincount(idx.clamp_(0, len(edges) - 1), minlength=len(edges)).double().cpu().numpy()
    cdf = 2.0 * norm.cdf(np.asarray(edges)) - 1.0
    p = np.diff(np.append(cdf, 1.0))
    expected = n * p
    return float(np.sum((counts - expected) ** 2 / expected)), counts, expected


def test_grid_normals_follow_the_normal_law(draws, record):
    torch, z, _ = draws
    n = z.numel()
    assert bool(torch.isfinite(z).all())
    assert bool((z.half().float() == z).all())                       # every value is an IEEE-half number
    assert float(z.abs().max()) < 6.8                                   # sqrt(-2 ln 2^-33)
    # cells of width 1/16 up to 5 sigma; an edge g > 0 is moved to the midpoint above the grid point g, so
    # that "rounded value > g" <=> "unrounded value > edge"
    grid = [i / 16.0 for i in range(0, 81)]
    edges = [0.0] + [g + _half_ulp(g) for g in grid[1:]]
    stat, counts, expected = _chi_square(torch, z.abs(), edges)
    dof = len(edges) - 1
    z64 = z.double()
    m1, m2 = float(z64.mean()), float((z64 * z64).mean())
    m3, m4 = float((z64 ** 3).mean()), float((z64 ** 4).mean())
    pos, neg, zero = int((z > 0).sum()), int((z < 0).sum()), int((z == 0).sum())
    t4, t5 = int((z.abs() > 4.0 + _half_ulp(4.0)).sum()), int((z.abs() > 5.0 + _half_ulp(5.0)).sum())
    e4 = n * 2 * norm.sf(4.0 + _half_ulp(4.0))
    e5 = n * 2 * norm.sf(5.0 + _half_ulp(5.0))
    record("philox_grid_law", chi2=stat, dof=dof, mean=m1, var=m2, third=m3, fourth=m4, pos=pos, neg=neg,
           zeros=zero, tail4=t4, tail4_expected=e4, tail5=t5, tail5_expected=e5)
    assert stat < dof + 6.0 * math.sqrt(2.0 * dof), (stat, dof)
    assert abs(m1) < 6.0 / math.sqrt(n) and abs(m2 - 1.0) < 6.0 * math.sqrt(2.0 / n)
    assert abs(m3) < 6.0 * math.sqrt(15.0 / n) and abs(m4 - 3.0) < 6.0 * math.sqrt(96.0 / n)
    assert abs(pos - neg) < 6.0 * math.sqrt(n) and zero < 64
    assert abs(t4 - e4) < 6.0 * math.sqrt(e4) and abs(t5 - e5) < 6.0 * math.sqrt(e5) + 1


def test_subnormal_region_near_zero(draws, record):
    """Half precision is subnormal below 2^-14 (spacing 2^-24): the normals that land there must keep their
    magnitude (no flush to zero) and arrive at the rate Phi predicts."""
    torch, z, _ = draws
    n = z.numel()
    lim = 2.0 ** -14
    small = z[z.abs() < lim]
    expected = n * (2.0 * norm.cdf(lim) - 1.0)
    k = small.numel()
    nz = int((small != 0).sum())
    distinct = int(torch.unique(small).numel())
    on_grid = bool(((small.double() * 2.0 ** 24).round() == small.double() * 2.0 ** 24).all())
    record("philox_subnormals", count=k, expected=expected, nonzero=nz, distinct=distinct)
    assert abs(k - expected) < 6.0 * math.sqrt(expected)
    assert on_grid and nz > 0.95 * k and distinct > 0.3 * k


def test_unrounded_normals_and_rounding_relation(draws, record):
    torch, z, zf = draws
    n = z.numel()
    assert bool(torch.isfinite(zf).all())
    # the grid value is the nearest half of the unrounded one
    assert bool((zf.half().float() == z).all())
    rel = float(((zf - z).abs() / zf.abs().clamp_min(2.0 ** -14)).max())
    edges = [i / 16.0 for i in range(0, 81)]
    stat, _, _ = _chi_square(torch, zf.abs(), edges)
    dof = len(edges) - 1
    var = float((zf.double() ** 2).mean())
    record("philox_full_law", chi2=stat, dof=dof, var=var, max_rel_rounding=rel)
    assert rel <= 2.0 ** -11 * 1.0001
    assert stat < dof + 6.0 * math.sqrt(2.0 * dof)
    assert abs(var - 1.0) < 6.0 * math.sqrt(2.0 / n)
    # the variance the rounding adds: ulp^2 / 12 relative, far below the Monte Carlo error of any moment match
    infl = float((z.double() ** 2).mean()) / var - 1.0
    assert abs(infl) < 1e-6


def test_columns_and_rows_are_uncorrelated(draws):
    """Lag-1 correlations along both axes of the (rows, cols) matrix and between the two normals of a
    Box-Muller pair."""
    torch, z, _ = draws
    m = z.reshape(ROWS, COLS).double()
    n = m.numel()
    c_col = float((m[:, 1:] * m[:, :-1]).mean())
    c_row = float((m[1:, :] * m[:-1, :]).mean())
    c_pair = float((m[:, 0::2] * m[:, 1::2]).mean())
    for c in (c_col, c_row, c_pair):
        assert abs(c) < 6.0 / math.sqrt(n / 2)
