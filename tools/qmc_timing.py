#!/usr/bin/env python
"""Throughput of the device-side scrambled-Sobol + erfinv generator (ppi_sobol_normal)
against SciPy's engine + scipy.special.erfinv on the host (what ppi/samplers.py:31-37 runs).

    python tools/qmc_timing.py [--rows 1048576] [--cols 2048]
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
    ap.add_argument("--rows", type=int, default=1 << 20)
    ap.add_argument("--cols", type=int, default=2048)
    ap.add_argument("--host-rows", type=int, default=1 << 13)
    a = ap.parse_args()
    from scipy.special import erfinv
    from scipy.stats import qmc
    from ppi_b200 import ops, samplers as smp

    s = smp.QuasiMonteCarlo(a.cols, seed=1)
    sv, sh = s._device_state()
    ops.sobol_normal(1024, a.cols, sv, sh, s.bits, 0)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    times = []
    for it in range(3):
        e0.record()
        z = ops.sobol_normal(a.rows, a.cols, sv, sh, s.bits, it * a.rows)
        e1.record()
        torch.cuda.synchronize()
        times.append(e0.elapsed_time(e1))
        del z
    ms = float(np.median(times))
    eng = qmc.Sobol(d=a.cols, scramble=True, seed=1)
    t0 = time.perf_counter()
    ref = np.sqrt(2) * erfinv(2 * 0.9999 * eng.random(a.host_rows) - 1)
    host_s = time.perf_counter() - t0
    n = a.rows * a.cols
    print(json.dumps({"rows": a.rows, "cols": a.cols, "device_ms": ms, "device_normals_per_s": n / ms * 1e3,
                      "device_write_GBps": 4.0 * n / ms * 1e-6,
                      "host_rows": a.host_rows, "host_normals_per_s": ref.size / host_s,
                      "speedup": (n / ms * 1e3) / (ref.size / host_s)}))


if __name__ == "__main__":
    main()
