#!/usr/bin/env python
"""Smallest possible exercise of the tensor-core sampler (run under `timeout` first after touching its
pipeline): one keep draw at (K, D), compared with mu + z L^T from the column-quad matrix."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "monte-carlo-posterior-policy-iteration_b200")):
    sys.path.insert(0, p)
import numpy as np
import torch

from ppi_b200 import ops

for K, D in ((512, 256), (3000, 384), (4096, 2048)):
    rng = np.random.default_rng(K)
    a = rng.standard_normal((D, D + 8))
    L = np.linalg.cholesky(a @ a.T / D + 0.2 * np.eye(D)).astype(np.float32)
    mu = rng.standard_normal(D).astype(np.float32)
    x, base = ops.sample_gaussian_keep(K, mu, L, seed=3, draw_id=1, row_offset=0)
    torch.cuda.synchronize()
    z = ops.philox_normal(K, D, 3, 1, 0, order="col").double().cpu().numpy()
    ref = mu[None, :].astype(np.float64) + z @ L.astype(np.float64).T
    err = np.max(np.abs(x.double().cpu().numpy() - ref)) / np.max(np.abs(ref))
    x2, b2, c = ops.sample_gaussian_cost(K, mu, L, "Rastrigin", seed=3, draw_id=1, row_offset=0, want_x=True)
    torch.cuda.synchronize()
    c_ref = 10.0 * D + (ref * ref - 10.0 * np.cos(2 * np.pi * ref)).sum(1)
    cerr = np.max(np.abs(c.double().cpu().numpy() - c_ref)) / np.max(np.abs(c_ref))
    print(f"K={K} D={D} x err {err:.2e} equal {bool(torch.equal(x, x2))} cost err {cerr:.2e}", flush=True)
print("sampler smoke ok")
