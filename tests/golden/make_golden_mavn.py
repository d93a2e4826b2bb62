"""
Golden fixtures of SURVEY 8(f) row 4: the module-level moment matchers of the reference called directly
(`m_projection`, `m_projection_mavn` with `iterations > 1` and the `update_out=True` flip-flop,
ppi/policies.py:78-143) and the matrix-normal KL / `vec` helpers (:161-210), from the UNMODIFIED reference.

    python tests/golden/make_golden_mavn.py      # build container only
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from make_golden import OUT, f32, import_reference  # noqa: E402


def main():
    alg, pol, smp, fn = import_reference()
    rng = np.random.default_rng(21)
    g = {}
    K, H, d = 200, 9, 4
    a = rng.standard_normal((H, H + 2))
    l_in = np.linalg.cholesky(a @ a.T / H + 0.2 * np.eye(H))
    l_out = np.diag([0.4, 0.8, 1.1, 0.6]) + 0.1 * np.tril(rng.standard_normal((d, d)), -1)
    xs = f32(0.3 + np.einsum("ki,bij,lj->bkl", l_in, rng.standard_normal((K, H, d)), l_out))
    log_w = f32(-2.5 * rng.random(K))
    cov_in0, cov_out0 = np.eye(H), np.diag([0.5, 1.0, 2.0, 1.5])
    g["xs"], g["log_w"], g["cov_in0"], g["cov_out0"] = xs, log_w, cov_in0, cov_out0
    mu, sigma, ess = pol.m_projection(log_w, xs.reshape(K, -1))
    g["mp_mu"], g["mp_sigma"], g["mp_ess"] = mu, sigma, ess
    for iters, upd in ((1, False), (1, True), (3, True), (4, False)):
        mean, cin, cout, ess = pol.m_projection_mavn(log_w, xs, cov_in0, cov_out0, iterations=iters,
                                                     update_out=upd)
        tag = f"mavn_{iters}_{int(upd)}"
        g[tag + "_mean"], g[tag + "_cin"], g[tag + "_cout"], g[tag + "_ess"] = mean, cin, cout, ess
    m1, m2 = rng.standard_normal((H, d)), rng.standard_normal((H, d))
    c1, c2 = g["mavn_3_1_cin"], cov_in0 + 0.1
    o1, o2 = g["mavn_3_1_cout"], cov_out0
    g["kl_m1"], g["kl_m2"], g["kl_c2"] = m1, m2, c2
    g["kl"] = pol.matrix_gaussian_kl(m1, c1, o1, m2, c2, o2)
    g["kl_same"] = pol.matrix_gaussian_kl(m1, c1, o1, m1, c1, o1)
    g["vec2"], g["vec3"] = pol.vec(m1), pol.vec(xs[:3].astype(np.float64))
    np.savez_compressed(os.path.join(OUT, "reference_golden_mavn.npz"), **g)
    print("wrote", len(g), "arrays; kl", g["kl"], g["kl_same"])


if __name__ == "__main__":
    main()
