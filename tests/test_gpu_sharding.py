"""Multi-rank correctness on real GPUs: the same seeded updates computed by one rank alone and
by two torchrun ranks (Philox rows sharded, costs all-gathered, moments all-reduced over NCCL, elite rows
fetched from their owners) agree to fp32 summation order; indices are identical.  Needs two visible GPUs
(`gpurun --gpus 2 -- python -m pytest tests/test_gpu_sharding.py -m gpu`); skipped on a one-GPU box."""
import json
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_two_rank_updates_equal_one_rank():
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    port = str(29600 + os.getpid() % 300)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", port, os.path.join(ROOT, "tools", "check_sharding.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    line = [ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1]
    out = json.loads(line)
    assert out["world"] == 2
    for kind in ("gaussian", "kernel", "icem"):
        assert all(v < 1e-5 for v in out[kind].values()), out
