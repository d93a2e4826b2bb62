#!/usr/bin/env python
"""Diagnostic: device time of ppi_solve_temperature (ESS target) at K = 2^20 .. 2^23."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "monte-carlo-posterior-policy-iteration_b200")):
    sys.path.insert(0, p)
import torch

from ppi_b200 import _lib, ops

for logk in (12, 16, 20, 23):
    K = 1 << logk
    c = torch.randn(K, device="cuda") * 30 + 1000
    for _ in range(3):
        ops.solve_temperature(c, _lib.TEMP_ESS_TARGET, K / 64, sync=False)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        nw, lw, st = ops.solve_temperature(c, _lib.TEMP_ESS_TARGET, K / 64, sync=False)
    e1.record()
    torch.cuda.synchronize()
    from ppi_b200.device import read_stats
    s = read_stats(st)
    print(f"K=2^{logk}: {e0.elapsed_time(e1) / 10 * 1e3:.1f} us per solve, {s.iterations} evaluations, alpha {s.alpha:.4f}")
