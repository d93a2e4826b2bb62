#!/usr/bin/env python
"""Latency of the MPC time shift per control step at C3 shapes (H = 30, d = 28):
the one-launch fp64 device kernel (ppi_kernel_shift) against the three-Gram-launch + NumPy host path it
replaced and against the oracle's pure NumPy restatement of the reference (ppi/policies.py:761-791).

    python tools/shift_timing.py [--horizon 30] [--steps 200]
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
    ap.add_argument("--horizon", type=int, default=30)
    ap.add_argument("--steps", type=int, default=200)
    a = ap.parse_args()
    from oracle import ppi_oracle as O
    from ppi_b200 import policies as pol, samplers as smp

    H, d, dt = a.horizon, 28, 0.01
    lo, hi = -np.ones(d), np.ones(d)

    def fresh():
        mean, cin, cout = pol.design_moments(hi, lo, ratio=1000.0)
        p = pol.Matern32Kernel(time_sequence=dt * np.arange(H), action_dimension=d, mean=mean, covariance_in=cin,
                               covariance_out=cout, lengthscale=0.05, sampler=smp.MonteCarlo,
                               limiter=pol.Limiter(lower=lo, upper=hi))
        p.compute_prior(p.t)
        p.covariance_in = 0.5 * p.covariance_in + 0.01 * np.eye(H)
        p.mean = 0.1 * np.random.default_rng(0).standard_normal((H, d))
        return p

    def run(p, steps):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for i in range(1, steps + 1):
            p.update_timesteps(dt * np.arange(i, i + H), 1.0)
        torch.cuda.synchronize()
        return (time.perf_counter() - t0) / steps * 1e6

    dev = fresh()
    run(dev, 10)
    dev = fresh()
    us_dev = run(dev, a.steps)
    # the kernel alone (CUDA events around the launch)
    from ppi_b200 import _lib
    real_call, spans = _lib.call, []

    def timed_call(name, *args):
        if name != "ppi_kernel_shift":
            return real_call(name, *args)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        rc = real_call(name, *args)
        e1.record()
        spans.append((e0, e1))
        return rc

    _lib.call = timed_call
    run(fresh(), 50)
    _lib.call = real_call
    us_kernel = float(np.median([e0.elapsed_time(e1) for e0, e1 in spans])) * 1e3
    host = fresh()
    host.kernel_kind = None          # the previous path: Gram launches + NumPy on the host mirrors
    run(host, 10)
    host = fresh()
    host.kernel_kind = None
    us_host = run(host, a.steps)
    err = float(np.max(np.abs(dev.covariance_in - host.covariance_in)) / np.max(np.abs(host.covariance_in)))

    # the reference's NumPy path (oracle restatement), Gram matrices on the host
    ref = fresh()
    gf = lambda t1, t2: O.gram("matern32", t1, t2, ref.sigma, ls=ref.ls)  # noqa: E731
    prior = gf(ref.t, ref.t)
    pinv = np.linalg.inv(prior)
    mean, cov, t = ref.mean, ref.covariance_in, ref.t
    t0 = time.perf_counter()
    for i in range(1, a.steps + 1):
        tn = dt * np.arange(i, i + H)
        mean, cov, _ = O.kernel_shift(gf, t, tn, prior, pinv, cov, mean, ref.mean_fn, lo, hi, 1.0, ref.sigma)
        t = tn
    us_ref = (time.perf_counter() - t0) / a.steps * 1e6
    print(json.dumps({"horizon": H, "d": d, "steps": a.steps, "us_per_shift_device_kernel": us_dev, "us_kernel_alone": us_kernel,
                      "us_per_shift_host_mirror": us_host, "us_per_shift_numpy_reference": us_ref,
                      "device_vs_host_cov_err": err}))


if __name__ == "__main__":
    main()
