#!/usr/bin/env python
"""Diagnostic: cProfile of the resident bench step of a workload (host-side overhead of small configs).
python tools/step_profile.py C3"""
import cProfile
import os
import pstats
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "monte-carlo-posterior-policy-iteration_b200")):
    sys.path.insert(0, p)
import torch

import bench

wl = dict(bench.WORKLOADS[sys.argv[1] if len(sys.argv) > 1 else "C3"])
policy, solver, cost, K = bench.build_workload(wl, 1)
for _ in range(5):
    bench.one_step(policy, solver, cost, K)
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
for _ in range(50):
    bench.one_step(policy, solver, cost, K)
torch.cuda.synchronize()
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(28)
