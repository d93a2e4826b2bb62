"""GPU tests of the hand-written dense D x D fp64 tail (csrc/dense64.cu): blocked Cholesky
on fp64 tensor cores, the triangular products L G and L L^T, and the fused whitened finish -- against
NumPy / LAPACK in fp64 (reference: np.linalg.cholesky at ppi/policies.py:261, det / solve of :146-158,213-216).
fp64 throughout, so the bar is 1e-11 relative, not the fp32 path's 1e-5."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def nerr(a, ref):
    a, ref = np.asarray(a, np.float64), np.asarray(ref, np.float64)
    assert a.shape == ref.shape, (a.shape, ref.shape)
    return np.max(np.abs(a - ref)) / max(np.max(np.abs(ref)), 1e-300)


@pytest.fixture(scope="module")
def ops():
    import torch

    assert torch.cuda.is_available()
    from ppi_b200 import ops as _ops

    return _ops


def _spd(D, rng, cond_shift=0.2):
    a = rng.standard_normal((D, D + 16))
    return a @ a.T / D + cond_shift * np.eye(D)


@pytest.mark.parametrize("D", [1, 63, 64, 65, 130, 200, 256, 1000, 1024, 2048])
def test_cholesky_blocked_matches_lapack(ops, record, D):
    rng = np.random.default_rng(D)
    a = _spd(D, rng)
    l64, l32, scal = ops.cholesky_blocked(a, want32=True)
    ref = np.linalg.cholesky(a)
    status, logdet = scal.cpu().numpy()[:2]
    got = l64.cpu().numpy()
    e = nerr(got, ref)
    record(f"cholesky_blocked[{D}]", chol=e, logdet=abs(logdet - 2 * np.log(np.diag(ref)).sum()))
    assert status == 0.0 and e < 1e-11
    assert np.array_equal(got, np.tril(got))                                   # zero above the diagonal
    assert abs(logdet - 2.0 * np.log(np.diag(ref)).sum()) < 1e-9 * max(1.0, abs(logdet))
    assert nerr(l32.cpu().numpy(), ref.astype(np.float32)) < 1e-6


@pytest.mark.parametrize("D,bad", [(200, 137), (1024, 0), (1024, 700), (300, 299)])
def test_cholesky_blocked_reports_the_failing_pivot(ops, D, bad):
    """A matrix that is positive definite up to pivot `bad` and indefinite there: LAPACK's potrf reports
    info = bad + 1 (NumPy raises LinAlgError); so does the blocked kernel, and NaNs fail too."""
    import scipy.linalg

    rng = np.random.default_rng(D + bad)
    a = _spd(D, rng)
    a[bad, bad] = -1.0
    _, info = scipy.linalg.lapack.dpotrf(a, lower=1)
    _, _, scal = ops.cholesky_blocked(a)
    assert int(scal[0].item()) == info == bad + 1
    b = _spd(D, rng)
    b[bad, bad // 2] = b[bad // 2, bad] = np.nan
    _, _, scal = ops.cholesky_blocked(b)
    assert scal[0].item() > 0


@pytest.mark.parametrize("D", [130, 384, 1000, 2048])
def test_tri_gram(ops, D):
    rng = np.random.default_rng(D + 1)
    L = np.linalg.cholesky(_spd(D, rng))
    out64, out32 = ops.tri_gram(ops.f64(L), want32=True)
    ref = L @ L.T
    got = out64.cpu().numpy()
    assert nerr(got, ref) < 1e-13 and np.array_equal(got, got.T)
    assert nerr(out32.cpu().numpy(), ref) < 1e-6


@pytest.mark.parametrize("D", [130, 200, 640, 1024, 2048])
def test_white_finish_against_numpy(ops, record, D):
    """C = S - m m^T, G = chol(C), L' = L G, mu' = mu + L m, log det C, tr C, |m|^2 in one call."""
    import torch

    rng = np.random.default_rng(D + 2)
    L = np.linalg.cholesky(_spd(D, rng))
    mu = rng.standard_normal(D).astype(np.float32)
    m = 0.3 * rng.standard_normal(D)
    S = (_spd(D, rng, 0.5) + np.outer(m, m)).astype(np.float32)
    S = 0.5 * (S + S.T)
    l64, l32, mu_new, scal = ops.white_finish(torch.as_tensor(S, device="cuda"), ops.f64(m), ops.f64(L),
                                              torch.as_tensor(mu, device="cuda"))
    status, logdet, trace, msq = scal.cpu().numpy()
    C = S.astype(np.float64) - np.outer(m, m)
    G = np.linalg.cholesky(C)
    e = dict(l_new=nerr(l64.cpu().numpy(), L @ G), mu=nerr(mu_new.cpu().numpy(), mu + L @ m),
             logdet=abs(logdet - 2 * np.log(np.diag(G)).sum()), trace=abs(trace - np.trace(C)) / np.trace(C),
             msq=abs(msq - m @ m) / (m @ m))
    record(f"white_finish[{D}]", **e)
    assert status == 0.0
    assert e["l_new"] < 1e-11 and e["mu"] < 1e-6 and e["logdet"] < 1e-9 * max(1.0, abs(logdet))
    assert e["trace"] < 1e-12 and e["msq"] < 1e-12
    got = l64.cpu().numpy()
    assert np.array_equal(got, np.tril(got))
    assert nerr(l32.cpu().numpy(), (L @ G).astype(np.float32)) < 1e-6
    # a singular C (rank one short) fails and leaves the outputs untouched
    S2 = S.copy()
    v = np.linalg.eigh(C)[1][:, 0]
    lam = np.linalg.eigh(C)[0][0]
    S2 = (S.astype(np.float64) - 1.5 * lam * np.outer(v, v)).astype(np.float32)
    _, _, _, scal2 = ops.white_finish(torch.as_tensor(S2, device="cuda"), ops.f64(m), ops.f64(L),
                                      torch.as_tensor(mu, device="cuda"))
    assert scal2[0].item() > 0 and abs(scal2[3].item() - m @ m) < 1e-12 * (m @ m)


@pytest.mark.parametrize("D", [1, 63, 64, 65, 130, 200, 840, 1000, 2048])
def test_tri_solve_sumsq_against_scipy(ops, record, D):
    """The solve-dependent KL terms (ppi/policies.py:146-158) from two Cholesky factors: |L^-1 B|_F^2 and
    |L^-1 v|^2 against SciPy's triangular solve in fp64; either operand alone as well."""
    import scipy.linalg
    import torch

    rng = np.random.default_rng(7 * D + 1)
    l = np.linalg.cholesky(_spd(D, rng))
    b = np.linalg.cholesky(_spd(D, rng, 0.5))
    v = rng.standard_normal(D)
    ref_b = float(np.sum(scipy.linalg.solve_triangular(l, b, lower=True) ** 2))
    ref_v = float(np.sum(scipy.linalg.solve_triangular(l, v, lower=True) ** 2))
    dev = torch.device("cuda")
    l_d, b_d, v_d = (torch.as_tensor(x, device=dev).contiguous() for x in (l, b, v))
    out = ops.tri_solve_sumsq(l_d, b_d, v_d).cpu().numpy()
    e_b, e_v = abs(out[0] - ref_b) / ref_b, abs(out[1] - ref_v) / ref_v
    record(f"tri_solve_sumsq[{D}]", matrix=e_b, vector=e_v)
    assert e_b < 1e-11 and e_v < 1e-11
    only_b = ops.tri_solve_sumsq(l_d, b_d, None).cpu().numpy()
    only_v = ops.tri_solve_sumsq(l_d, None, v_d).cpu().numpy()
    # (per-CTA partial sums meet in an fp64 atomic: equal up to summation order)
    assert abs(only_b[0] - out[0]) <= 1e-13 * out[0] and only_b[1] == 0.0
    assert abs(only_v[1] - out[1]) <= 1e-13 * out[1] and only_v[0] == 0.0
    # identity right-hand side: |L^-1|_F^2 (the regularisation branch of the whitened update)
    eye = torch.eye(D, dtype=torch.float64, device=dev)
    inv = float(ops.tri_solve_sumsq(l_d, eye, None)[0].item())
    ref_inv = float(np.sum(scipy.linalg.solve_triangular(l, np.eye(D), lower=True) ** 2))
    assert abs(inv - ref_inv) / ref_inv < 1e-11
