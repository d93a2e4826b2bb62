import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "monte-carlo-posterior-policy-iteration_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def golden():
    import numpy as np

    return dict(np.load(os.path.join(ROOT, "tests", "golden", "reference_golden.npz")))


def record_errors(name, **values):
    """Append the ACHIEVED errors of a parity test to gpurun_out/parity_errors.jsonl (and print them), so that
    the measured margins -- not only pass / fail -- come home from the GPU box (profiles/ keeps a copy)."""
    import json

    line = {"test": name}
    line.update({k: (float(v) if isinstance(v, (int, float)) or hasattr(v, "__float__") else v)
                 for k, v in values.items()})
    print("PARITY", json.dumps(line))
    out = os.path.join(ROOT, "gpurun_out")
    try:
        os.makedirs(out, exist_ok=True)
        with open(os.path.join(out, "parity_errors.jsonl"), "a") as f:
            f.write(json.dumps(line) + "\n")
    except OSError:
        pass


@pytest.fixture(scope="session")
def record():
    return record_errors
