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


K = int(sys.argv[1]) if len(sys.argv) > 1 else wl["K"] // 8      # the per-rank share at N = 8


def step():
    pol_, sol_, cost_ = bench.build_workload(wl, K, pinned=host)
    bench.one_step(pol_, sol_, cost_, K)
    sink["mu"], sink["sigma"] = pol_.mu, pol_.sigma


for _ in range(3):
    step()
torch.cuda.synchronize()
import time

t0 = time.perf_counter()
for _ in range(10):
    step()
torch.cuda.synchronize()
print(f"e2e step at K = {K}: {(time.perf_counter() - t0) * 100:.3f} ms wall per step")
pr = cProfile.Profile()
pr.enable()
for _ in range(5):
    step()
torch.cuda.synchronize()
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(45)
