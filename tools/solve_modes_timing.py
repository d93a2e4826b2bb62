import os, sys
ROOT = "/root/repo"
for p in (ROOT, os.path.join(ROOT, "monte-carlo-posterior-policy-iteration_b200")):
    sys.path.insert(0, p)
import torch
from ppi_b200 import _lib, ops
from ppi_b200.device import read_stats
for name, mode, hyp in (("ess", _lib.TEMP_ESS_TARGET, None), ("lbps", _lib.TEMP_LOWER_BOUND, 0.9), ("reps", _lib.TEMP_REPS_KL, 1.0)):
    for logk in (10, 12, 16, 20):
        K = 1 << logk
        c = torch.randn(K, device="cuda") * 30 + 1000
        h = (K / 64 if K >= 4096 else 10) if hyp is None else hyp
        for _ in range(3):
            ops.solve_temperature(c, mode, h, sync=False)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            nw, lw, st = ops.solve_temperature(c, mode, h, sync=False)
        e1.record()
        torch.cuda.synchronize()
        s = read_stats(st)
        print(f"{name} K=2^{logk}: {e0.elapsed_time(e1) / 10 * 1e3:.1f} us per solve, {s.iterations} evaluations, alpha {s.alpha:.4f}, {e0.elapsed_time(e1) / 10 * 1e3/ max(1,s.iterations):.2f} us/eval")
