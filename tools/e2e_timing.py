#!/usr/bin/env python
"""Diagnostic: where the host side of bench.py's e2e leg (fresh policy from pinned host state every
step, results read back) spends its time.  python tools/e2e_timing.py"""
import cProfile
import os
import pstats
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "monte-carlo-posterior-policy-iteration_b200")):
    sys.path.insert(0, p)
import torch

import bench

wl = dict(bench.WORKLOADS["C5"])
D = wl["D"]
pin = {"mu": torch.ones(D, dtype=torch.float64).pin_memory(),
       "sigma": (0.5 * torch.eye(D, dtype=torch.float64)).pin_memory()}
host = {k: v.numpy() for k, v in pin.items()}
sink = {}


def step():
    pol_, sol_, cost_, K = bench.build_workload(wl, 1, pinned=host)
    bench.one_step(pol_, sol_, cost_, K)
    sink["mu"], sink["sigma"] = pol_.mu, pol_.sigma


for _ in range(3):
    step()
torch.cuda.synchronize()
pr = cProfile.Profile()
pr.enable()
for _ in range(5):
    step()
torch.cuda.synchronize()
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(35)
