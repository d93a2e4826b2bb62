#!/usr/bin/env python
"""K-sharding check on real GPUs: the same seeded update computed by one rank alone
and by all ranks of a torchrun job (Philox rows sharded, costs all-gathered, whitened moments
all-reduced over NCCL) must agree to fp32 summation order.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
      --master-port 29533 tools/check_sharding.py
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "monte-carlo-posterior-policy-iteration_b200")):
    sys.path.insert(0, p)
import numpy as np
import torch
import torch.distributed as dist


def run(kind, K, iters=2):
    from ppi_b200 import algorithms as alg, functions as fn, policies as pol, samplers as smp

    if kind == "icem":
        # iCEM + Particles on a ColouredNoise prior: the re-used elites are GLOBAL sample numbers whose rows
        # live on either rank (ppi/algorithms.py:162-164)
        H, d = 12, 4
        t = 0.05 * np.arange(H)
        np.random.seed(5)
        policy = pol.ColouredNoise(time_sequence=t, action_dimension=d, mean=np.zeros(d),
                                   covariance_in=np.array([1.0]), covariance_out=0.3 * np.eye(d),
                                   sampler=smp.Particles, beta=1.5,
                                   limiter=pol.Limiter(lower=-np.ones(d), upper=np.ones(d)))
        policy._rng = smp.MonteCarlo(H * d, seed=99)
        solver, cost = alg.iCem(n_elites=64), fn.TrackingCost(t, d)
        solver.reset(policy)
        for _ in range(3):
            a, p = policy(K)
            solver.update(cost(a), p, policy)
        return policy.mean.copy(), policy.std.copy(), np.asarray(policy.sampler.particles).copy()
    if kind == "gaussian":
        D = 512
        policy = pol.GaussianPolicy(np.ones(D), 0.5 * np.eye(D), smp.MonteCarlo(D, seed=99))
        solver, cost = alg.Essps(n_elites=K // 32), fn.Rastrigin(D)
    else:
        H, d = 30, 28
        t = 0.01 * np.arange(H)
        mean, cin, cout = pol.design_moments(np.ones(d), -np.ones(d), ratio=1000.0)
        policy = pol.SquaredExponentialKernel(
            time_sequence=t, action_dimension=d, mean=mean, covariance_in=cin, covariance_out=cout,
            lengthscale=0.05, sampler=lambda dim: smp.MonteCarlo(dim, seed=99),
            limiter=pol.Limiter(lower=-np.ones(d), upper=np.ones(d)))
        solver, cost = alg.Lbps(delta=0.9), fn.TrackingCost(t, d)
    solver.reset(policy)
    for _ in range(iters):
        a, p = policy(K)
        solver.update(cost(a), p, policy)
    if kind == "gaussian":
        return policy.mu.copy(), policy.sigma.copy(), float(solver.trace["alpha"][-1])
    return policy.mean.copy(), policy.covariance_in.copy(), float(solver.trace["alpha"][-1])


def main():
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    K = 1 << 17
    kinds = ("gaussian", "kernel", "icem")
    single = {k: run(k, K) for k in kinds}          # before init: world size 1
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    sharded = {k: run(k, K) for k in kinds}
    out = {"world": dist.get_world_size()}
    for k in single:
        (m1, s1, a1), (m2, s2, a2) = single[k], sharded[k]
        a1, a2 = np.asarray(a1, np.float64), np.asarray(a2, np.float64)
        out[k] = {"mean_err": float(np.max(np.abs(m1 - m2)) / np.max(np.abs(m1))),
                  "cov_err": float(np.max(np.abs(s1 - s2)) / np.max(np.abs(s1))),
                  "alpha_err": float(np.max(np.abs(a1 - a2)) / np.max(np.abs(a1)))}   # icem: the particles
    if dist.get_rank() == 0:
        print(json.dumps(out))
        assert all(v < 1e-5 for k in single for v in out[k].values()), out
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
