#!/usr/bin/env python
"""Diagnostic (not part of the product path): wall / device time of each phase of one C5-style
GaussianPolicy update, plus cuBLAS TF32 / BF16 matmul rates on this box for the roofline notes.

  python tools/phase_timing.py [--K 1048576] [--D 2048]
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "monte-carlo-posterior-policy-iteration_b200")):
    sys.path.insert(0, p)

import numpy as np
import torch


def sync_time(fn, n=3):
    torch.cuda.synchronize()
    out = []
    for _ in range(n):
        t0 = time.perf_counter()
        r = fn()
        torch.cuda.synchronize()
        out.append((time.perf_counter() - t0) * 1e3)
    return min(out), r


def matmul_peak(dtype, tf32, n=8192, reps=20):
    torch.backends.cuda.matmul.allow_tf32 = tf32
    a = torch.randn(n, n, device="cuda", dtype=dtype)
    b = torch.randn(n, n, device="cuda", dtype=dtype)
    for _ in range(3):
        a @ b
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        a @ b
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    t0 = time.perf_counter()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    m = 0
    while time.perf_counter() - t0 < 2.0:
        for _ in range(10):
            a @ b
        m += 10
        torch.cuda.synchronize()
    e1.record()
    torch.cuda.synchronize()
    sustained = e0.elapsed_time(e1) / m
    torch.backends.cuda.matmul.allow_tf32 = False
    return 2.0 * n ** 3 / best / 1e9, 2.0 * n ** 3 / sustained / 1e9


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--K", type=int, default=1 << 20)
    ap.add_argument("--D", type=int, default=2048)
    args = ap.parse_args()
    from ppi_b200 import _lib, algorithms as alg, functions as fn, ops, policies as pol, samplers as smp

    K, D = args.K, args.D
    res = {}
    res["tf32_tflops_burst_sustained"] = matmul_peak(torch.float32, True)
    res["bf16_tflops_burst_sustained"] = matmul_peak(torch.bfloat16, False)

    policy = pol.GaussianPolicy(np.ones(D), 0.5 * np.eye(D), smp.MonteCarlo(D, seed=1234))
    solver = alg.Essps(n_elites=max(1, K // 64))
    cost = fn.Rastrigin(D)
    solver.reset(policy)
    for _ in range(2):
        a, p = policy(K)
        c = cost(a)
        solver.update(c, p, policy)
    ph = {}
    ph["sample"], (a, p) = sync_time(lambda: policy(K))
    ph["cost"], c = sync_time(lambda: cost(a))
    ph["solve"], w = sync_time(lambda: solver._solve(c, p, _lib.TEMP_ESS_TARGET, solver.ess))
    x = p.tensor
    mu_d = policy._mu_d()
    ph["wmoments_diag"], s1 = sync_time(lambda: ops.wmoments_diag(x, w.nw, shift1=mu_d))
    mu_new, _ = ops.finalize_gaussian(s1=s1, shift=mu_d, want_sigma=False)
    ph["wmoments_dense"], cov = sync_time(lambda: ops.wmoments_dense(x, w.nw, shift=mu_new))
    ph["cov.double()"], c64 = sync_time(lambda: cov.double())
    ph["cholesky_ex fp64"], _ = sync_time(lambda: torch.linalg.cholesky_ex(c64))
    ph["cholesky_ex fp32"], _ = sync_time(lambda: torch.linalg.cholesky_ex(cov))
    l64 = torch.linalg.cholesky(c64)
    ph["solve_triangular fp64"], _ = sync_time(lambda: torch.linalg.solve_triangular(l64, l64, upper=False))
    ph["sigma d2h fp64"], _ = sync_time(lambda: c64.cpu())
    ph["weighted_update total"], _ = sync_time(lambda: policy.weighted_update(w, p))
    ph["entropy"], _ = sync_time(lambda: policy.entropy)

    def full():
        a_, p_ = policy(K)
        c_ = cost(a_)
        solver._stats = None
        solver.update(c_, p_, policy)

    ph["full step"], _ = sync_time(full)
    res["phases_ms"] = ph
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
