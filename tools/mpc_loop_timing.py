#!/usr/bin/env python
"""Latency of one MPC control step at door-v0 shapes (C3: H = 30, d = 28) through the reference-facing
API, with the synthetic tracking return in place of the MuJoCo rollouts (mpc/mpc.py:58-84):

    policy.update_timesteps(t, anneal)
    n_iters x [ actions, params = policy(K); costs = cost(actions); solver.update(costs, params, policy) ]
    action = policy.predict(only_mean=True)[0]

against the oracle's NumPy restatement of the same step (naive einsum forms as the reference writes them,
and BLAS forms), on this machine's host cores.

    python tools/mpc_loop_timing.py [--samples 1024] [--iters 2] [--steps 50]
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--samples", type=int, default=1024)
    ap.add_argument("--iters", type=int, default=2)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--cpu-steps", type=int, default=3)
    ap.add_argument("--profile", action="store_true", help="cProfile the device loop and skip the NumPy legs")
    a = ap.parse_args()
    from oracle import ppi_oracle as O
    from ppi_b200 import algorithms as alg, functions as fn, policies as pol, samplers as smp

    H, d, dt, K = 30, 28, 0.01, a.samples
    lo, hi = -np.ones(d), np.ones(d)
    mean, cin, cout = pol.design_moments(hi, lo, ratio=1000.0)
    t0 = dt * np.arange(H)
    policy = pol.Matern32Kernel(time_sequence=t0, action_dimension=d, mean=mean, covariance_in=cin,
                                covariance_out=cout, lengthscale=0.05, sampler=lambda dim: smp.MonteCarlo(dim, seed=3),
                                limiter=pol.Limiter(lower=lo, upper=hi))
    policy.compute_prior(t0)
    solver = alg.Lbps(delta=0.9)
    solver.reset(policy)
    cost = fn.TrackingCost(t0, d)

    def control_step(i):
        policy.update_timesteps(dt * np.arange(i, i + H), 1.0)
        for _ in range(a.iters):
            actions, params = policy(K)
            solver.update(cost(actions), params, policy)
        return policy.predict(only_mean=True)[0]

    for i in range(1, 6):
        control_step(i)
    torch.cuda.synchronize()
    t_start = time.perf_counter()
    for i in range(6, 6 + a.steps):
        act = control_step(i)
    torch.cuda.synchronize()
    ms_dev = (time.perf_counter() - t_start) / a.steps * 1e3
    if a.profile:
        import cProfile
        import io
        import pstats

        pr = cProfile.Profile()
        pr.enable()
        for i in range(6 + a.steps, 6 + 2 * a.steps):
            control_step(i)
        torch.cuda.synchronize()
        pr.disable()
        s = io.StringIO()
        pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(60)
        print(f"control step {ms_dev:.3f} ms (K = {K}, iters = {a.iters})")
        print(s.getvalue()[:12000])
        return

    # the same control step in NumPy (oracle restatement of the reference), fresh standard normals
    rng = np.random.default_rng(0)
    gf = lambda t1, t2: O.gram("matern32", t1, t2, policy.sigma, ls=policy.ls)  # noqa: E731
    target = O.tracking_target(t0, d)

    def cpu_step(state, i, naive):
        t, prior, pinv, m, c, l = state
        tn = dt * np.arange(i, i + H)
        m, c, l = O.kernel_shift(gf, t, tn, prior, pinv, c, m, mean, lo, hi, 1.0, policy.sigma)
        l_out = np.linalg.cholesky(cout)
        for _ in range(a.iters):
            z = rng.standard_normal((K, H, d))
            xs = O.sample_matnorm(z, mean, m, l, l_out, lo, hi, naive=naive)
            costs = O.synthetic_return(xs, target)
            log_w, _ = O.lbps_log_w(costs, 0.9)
            corrected = xs - mean[None, None, :]
            m, c, _, _ = O.m_projection_mavn(log_w, corrected, c, cout, naive=naive)
            l = np.linalg.cholesky(c)
        return (tn, prior, pinv, m, c, l)

    prior = gf(t0, t0)
    out = {}
    for naive in (False, True):
        state = (t0, prior, np.linalg.inv(prior), np.zeros((H, d)), prior.copy(), np.linalg.cholesky(prior))
        state = cpu_step(state, 1, naive)
        t_start = time.perf_counter()
        for i in range(2, 2 + a.cpu_steps):
            state = cpu_step(state, i, naive)
        out["naive" if naive else "blas"] = (time.perf_counter() - t_start) / a.cpu_steps * 1e3
    print(json.dumps({"shapes": f"H={H} d={d} K={K} iters={a.iters}", "control_step_ms_b200": ms_dev,
                      "control_step_ms_numpy_reference_forms": out["naive"],
                      "control_step_ms_numpy_blas_forms": out["blas"],
                      "cpu_threads": os.cpu_count(), "first_action": [float(v) for v in act[:3]]}))


if __name__ == "__main__":
    main()
