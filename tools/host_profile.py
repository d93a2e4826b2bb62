#!/usr/bin/env python
"""cProfile of the host side of one small-matrix update loop (C3 by default): where the Python / ctypes / sync
time of a latency-bound workload goes.   python tools/host_profile.py [--workload C3] [--iters 400]"""
import argparse
import cProfile
import io
import os
import pstats
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "monte-carlo-posterior-policy-iteration_b200")):
    sys.path.insert(0, p)
import torch

import bench


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="C3")
    ap.add_argument("--iters", type=int, default=400)
    args = ap.parse_args()
    wl = dict(bench.WORKLOADS[args.workload])
    policy, solver, cost = bench.build_workload(wl, wl["K"])
    for _ in range(20):
        bench.one_step(policy, solver, cost, wl["K"])
    policy, solver, cost = bench.build_workload(wl, wl["K"])
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    pr = cProfile.Profile()
    pr.enable()
    for i in range(args.iters):
        if i % 8 == 0:
            policy, solver, cost = bench.build_workload(wl, wl["K"])
        bench.one_step(policy, solver, cost, wl["K"])
    torch.cuda.synchronize()
    pr.disable()
    dt = time.perf_counter() - t0
    print(f"{args.workload}: {dt / args.iters * 1e3:.3f} ms per update under cProfile")
    s = io.StringIO()
    pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(45)
    print(s.getvalue()[:9000])


if __name__ == "__main__":
    main()
