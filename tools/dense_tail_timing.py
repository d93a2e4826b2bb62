#!/usr/bin/env python
"""Timing of the fp64 dense tail alone (csrc/dense64.cu): ppi_white_finish, ppi_cholesky_blocked, ppi_tri_gram at
D = 1024 / 2048, CUDA events around each call, next to torch.linalg (cuSOLVER / cuBLAS) for the same algebra.
    python tools/dense_tail_timing.py [--dims 1024 2048] [--reps 5]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "monte-carlo-posterior-policy-iteration_b200")):
    sys.path.insert(0, p)
import numpy as np
import torch


def timed(fn, reps):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--dims", type=int, nargs="+", default=[1024, 2048])
    ap.add_argument("--reps", type=int, default=5)
    args = ap.parse_args()
    from ppi_b200 import ops

    out = {}
    for D in args.dims:
        rng = np.random.default_rng(D)
        a = rng.standard_normal((D, D + 16))
        spd = a @ a.T / D + 0.3 * np.eye(D)
        L = torch.linalg.cholesky(torch.as_tensor(spd, device="cuda")).contiguous()
        m = torch.as_tensor(0.1 * rng.standard_normal(D), device="cuda")
        S = (torch.as_tensor(spd, device="cuda") + torch.outer(m, m)).float()
        mu = torch.zeros(D, dtype=torch.float32, device="cuda")
        a64 = torch.as_tensor(spd, device="cuda")

        def torch_tail():
            c = S.double()
            c.addr_(m, m, alpha=-1.0)
            g, info = torch.linalg.cholesky_ex(c)
            return L @ g, info

        out[D] = {
            "white_finish_ms": timed(lambda: ops.white_finish(S, m, L, mu), args.reps),
            "cholesky_blocked_ms": timed(lambda: ops.cholesky_blocked(a64, want32=True), args.reps),
            "tri_gram_ms": timed(lambda: ops.tri_gram(L), args.reps),
            "torch_cholesky_ex_plus_LG_ms": timed(torch_tail, args.reps),
            "torch_cholesky_ms": timed(lambda: torch.linalg.cholesky_ex(a64), args.reps),
        }
    # phase clocks of the critical-path CTAs of the last factorisation (debug export of dense64.cu)
    import ctypes as C

    from ppi_b200 import _lib

    lib = _lib.load()
    buf = (C.c_longlong * (64 * 2 * 8))()
    torch.cuda.synchronize()
    lib.ppi_debug_chol_trace.argtypes = [C.c_void_p]
    lib.ppi_debug_chol_trace(buf)
    tr = np.array(buf, dtype=np.int64).reshape(64, 2, 8)
    T = args.dims[-1] // 64
    ph = {}
    for cta, names in ((0, ("load", "update", "factor", None, "store_diag")), (1, ("load", "update", "factor", "solve", "store"))):
        d = np.diff(tr[1:T - 1, cta, :6].astype(np.float64), axis=1)      # steps 1 .. T-2: all phases present
        for k, nm in enumerate(names):
            if nm:
                ph[f"cta{cta}_{nm}_cycles"] = float(np.median(d[:, k]))
    ph["cta1_total_cycles"] = float(np.median(tr[1:T - 1, 1, 5] - tr[1:T - 1, 1, 0]))
    out["phase_clocks_last_dim"] = ph
    print(json.dumps(out))


if __name__ == "__main__":
    main()
