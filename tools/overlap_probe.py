"""Does the base-normal generator (ALU-bound) run UNDER the dense fp64 tail (latency-bound) when it is issued
on a side stream?  Times both alone and together, for several resident-CTA caps of the generator and with
the tail on a high-priority stream.   python tools/overlap_probe.py [K] [D]
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "monte-carlo-posterior-policy-iteration_b200"))
import numpy as np
import torch

from ppi_b200 import _lib, ops


def main():
    K = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
    D = int(sys.argv[2]) if len(sys.argv) > 2 else 2048
    dev = torch.device("cuda")
    rng = np.random.default_rng(0)
    a = rng.standard_normal((D, D + 64))
    c = a @ a.T / (D + 64)
    s = torch.as_tensor(c + np.outer(np.ones(D), np.ones(D)) * 1e-4, dtype=torch.float32, device=dev)
    m = torch.full((D,), 1e-2, dtype=torch.float64, device=dev)
    l_old = torch.linalg.cholesky(torch.as_tensor(c, dtype=torch.float64, device=dev)).contiguous()
    mu_old = torch.zeros(D, dtype=torch.float32, device=dev)
    sbytes = _lib.query("ppi_zstore_bytes", K, D, 0)
    store = ops._aligned_bytes(sbytes, dev)
    side = torch.cuda.Stream()
    hi = torch.cuda.Stream(priority=-1)

    def tail():
        return ops.white_finish(s, m, l_old, mu_old)

    def gen():
        return ops.philox_prefetch(K, D, 1, 1, 0, store, side, ("probe",))

    def timed(fn, stream=None, reps=5):
        stream = stream or torch.cuda.current_stream()
        out = []
        for _ in range(reps):
            torch.cuda.synchronize()
            with torch.cuda.stream(stream):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                fn()
                e1.record()
            torch.cuda.synchronize()
            out.append(e0.elapsed_time(e1))
        return float(np.median(out))

    tail()
    torch.cuda.synchronize()
    print(f"tail alone            {timed(tail):7.3f} ms")
    for blocks in (6, 2, 1):
        _lib.call("ppi_set_option", _lib.OPT_PREFETCH_BLOCKS, blocks)

        def gen_join():
            pre = gen()
            torch.cuda.current_stream().wait_event(pre.done)

        t_gen = timed(gen_join)
        for name, stream in (("default", None), ("high-priority", hi)):
            def both():
                pre = gen()
                tail()
                torch.cuda.current_stream().wait_event(pre.done)

            def tail_only_clock():
                # the tail's own duration while the generator runs beside it
                pre = gen()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                tail()
                e1.record()
                torch.cuda.current_stream().wait_event(pre.done)
                return e0, e1

            t_both = timed(both, stream)
            torch.cuda.synchronize()
            with torch.cuda.stream(stream or torch.cuda.current_stream()):
                e0, e1 = tail_only_clock()
            torch.cuda.synchronize()
            print(f"blocks {blocks}: generator alone {t_gen:6.3f} ms | together ({name} tail) {t_both:6.3f} ms, "
                  f"tail inside {e0.elapsed_time(e1):6.3f} ms")


if __name__ == "__main__":
    main()
