"""Diagnostic: error of the dense covariance vs the TMEM accumulation chain length (chunk_kb)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "monte-carlo-posterior-policy-iteration_b200")):
    sys.path.insert(0, p)
import numpy as np, torch
from ppi_b200 import _lib, ops

def run(K, D, seed, posmean):
    g = torch.Generator(device="cuda").manual_seed(seed)
    x = torch.randn(K, D, device="cuda", generator=g) * (0.2 + 3 * torch.rand(D, device="cuda", generator=g))
    if posmean:
        x = x + 3.0   # uncentred: every product positive-biased -> truncation bias would show
    lw = -8.0 * torch.rand(K, device="cuda", generator=g)
    nw = torch.softmax(lw.double(), 0).float()
    ref = ((nw.double()[:, None] * x.double()).T @ x.double())
    out = {}
    for scheme in (0, 1):
        _lib.call("ppi_set_option", 0, scheme)
        for chunk in (2, 4, 16, 64):
            _lib.call("ppi_set_option", 1, chunk)
            cov = ops.wmoments_dense(x, nw).double()
            out[(scheme, chunk)] = float((cov - ref).abs().max() / ref.abs().max())
    return out

for K, D, pm in ((65536, 256, False), (65536, 256, True), (262144, 256, True), (65536, 1024, True)):
    print(K, D, "uncentred" if pm else "centred", {k: f"{v:.2e}" for k, v in run(K, D, 1, pm).items()})
