#!/usr/bin/env python
"""bench.py -- PPI update samples/sec on B200 (BASELINE.json metric), one JSON line on stdout.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--workload C5|C2|C3|C4]

A "step" is one pass of the hot path (ppi/algorithms.py:85-88): sample K actions from the prior ->
cost -> Gibbs-posterior temperature + weights -> weighted moment match -> new prior.  Default
workload = config C5 of BASELINE.json: GaussianPolicy, D = 2048, K = 2^20 samples per GPU
(weak scaling: the K samples shard over the ranks, costs are all-gathered, moments all-reduced),
Rastrigin cost, ESSPS temperature with N = K/64, on-device Philox normals.

value      whole-job samples/s with the policy state resident in HBM (CUDA events, max over ranks)
e2e        same step through the public API with HOST buffers: a fresh policy is built from pinned
           host mu / sigma every step and the fitted mu / sigma are read back to the host
roofline   the dominant kernel (tcgen05 covariance GEMM): algorithmic flops per launch / its
           CUDA-event time measured inside the timed region, against the measured 16-bit tensor peak
cpu_baseline  the oracle port (NumPy restatement of the reference, its naive einsum) on a bounded
           sample of the same workload, timed on this box's host cores (rank 0, N = 1 only)
--impl reference  times that CPU path alone (the reference is pure NumPy: there is no GPU arm)
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "monte-carlo-posterior-policy-iteration_b200")
for _p in (ROOT, PKG):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402

WORKLOADS = {
    # name: (description, dims)
    "C5": dict(kind="gaussian", D=2048, K=2 ** 20, fn="Rastrigin",
               desc="C5: GaussianPolicy D=2048, K=2^20 samples per GPU, Rastrigin, Essps(N=K/64), Philox"),
    "C2": dict(kind="gaussian", D=1024, K=65536, fn="Rastrigin",
               desc="C2: GaussianPolicy D=1024, K=65536 per GPU, Rastrigin, Essps(N=K/64), Philox"),
    "C3": dict(kind="kernel", H=30, d=28, K=4096,
               desc="C3: SquaredExponentialKernel H=30 d=28 (door-v0 shapes), K=4096 per GPU, tracking return, Lbps(0.9)"),
    "C4": dict(kind="features", T=1000, d=2, K=65536,
               desc="C4: RffFeatures T=1000 F=21 d=2 (ball-in-a-cup shapes), K=65536 per GPU, tracking return, Essps(10)"),
}


# dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, from the committed
# `ncu --set full` capture under profiles/ (see profiles/README.md); keyed by (kernel, D).
# (kernel, D, samples per launch)
NCU_TRAFFIC = {
    ("syrk_pair", 2048, 65536): 541.2e6,     # profiles/r1_gemm_ncu_full_summary.csv (537.0 MB read + 4.2 MB written)
    ("syrk_white", 2048, 1 << 20): 4.3052e9,  # profiles/README.md: 4.301 GB read (= the 2-byte store, once) + 4.1 MB written
    ("syrk_white", 2048, 1 << 18): 1.0793e9,  # profiles/r1_syrk_white_ncu.csv
}


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return dict(hbm=p["hbm_gbs"], bf16_burst=p["bf16_tflops"], bf16=p["bf16_tflops_sustained"],
                    source="measured (MEASURED_PEAKS.json)")
    return dict(hbm=6650.0, bf16_burst=1590.0, bf16=1400.0, source="fallback (B200_PROFILING.md)")


# ------------------------------------------------------------------------------------------------
# clocks sampled during the timed region
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "20",
                 "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smax, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.lines:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                smax = float(f[1])
            except ValueError:
                continue
            for name, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": smax,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------------
# CPU arm: the oracle port of the reference
# ------------------------------------------------------------------------------------------------
def cpu_step(wl, K, rng, naive=True):
    """One update of the same workload with the NumPy oracle (reference restatement; naive=True keeps
    the reference's three-operand einsum, naive=False is the BLAS form of the same math)."""
    from oracle import ppi_oracle as O

    if wl["kind"] == "gaussian":
        D = wl["D"]
        mu, sigma = np.ones(D), 0.5 * np.eye(D)
        z = rng.standard_normal((K, D))
        x = O.sample_gaussian(z, mu, sigma)
        c = getattr(O, wl["fn"].lower())(x)
        log_w, _ = O.essps_log_w(c, max(1, K // 64))
        O.gaussian_weighted_update(mu, sigma, log_w, x, naive=naive)
    elif wl["kind"] == "kernel":
        H, d = wl["H"], wl["d"]
        t = 0.01 * np.arange(H)
        cin = O.gram("se", t, t, 1000.0, 0.05)
        l_in, cout = np.linalg.cholesky(cin), 1e-3 * np.eye(d)
        z = rng.standard_normal((K, H, d))
        x = O.sample_matnorm(z, np.zeros(d), np.zeros((H, d)), l_in, np.sqrt(cout), -np.ones(d), np.ones(d))
        c = O.synthetic_return(x, O.tracking_target(t, d))
        log_w, _ = O.lbps_log_w(c, 0.9)
        O.kernel_weighted_update(np.zeros(d), cin, l_in, cout, log_w, x)
    else:
        T, d = wl["T"], wl["d"]
        t = np.linspace(0, 2, T)
        feat, dfeat = O.rff_features(t, np.sqrt(3e-2), 10, True)
        F = feat.shape[1]
        cin, cout = 1e2 * np.eye(F), 1e-3 * np.eye(d)
        z = rng.standard_normal((K, F, d))
        ys, ws = O.sample_features(z, np.array([0.0, 1.5707]), np.zeros((F, d)), np.linalg.cholesky(cin),
                                   np.sqrt(cout), feat, dfeat)
        c = O.synthetic_return(ys, O.tracking_target(t, d))
        log_w, _ = O.essps_log_w(c, 10)
        O.features_weighted_update(np.zeros((F, d)), cin, np.linalg.cholesky(cin), cout, log_w, ws)


def cpu_sample_size(wl, target_s):
    """Rows of the bounded CPU sample so that one step takes about target_s seconds.  The per-row
    cost is timed on the K-proportional part alone (sampling + moment matching einsums); the
    K-independent part (O(D^3) factorisations) is timed once."""
    from oracle import ppi_oracle as O

    rng = np.random.default_rng(0)
    k = 32
    cpu_step(wl, 8, rng)  # warm caches / lazy imports
    t0 = time.perf_counter()
    cpu_step(wl, k, rng)
    t_step = time.perf_counter() - t0
    t0 = time.perf_counter()
    if wl["kind"] == "gaussian":
        O.m_projection(np.zeros(k), rng.standard_normal((k, wl["D"])))
    elif wl["kind"] == "kernel":
        x = rng.standard_normal((k, wl["H"], wl["d"]))
        O.sample_matnorm(x, np.zeros(wl["d"]), np.zeros((wl["H"], wl["d"])), np.eye(wl["H"]), np.eye(wl["d"]))
        O.m_projection_mavn(np.zeros(k), x, np.eye(wl["H"]), np.eye(wl["d"]))
    else:
        t_step, k = t_step, k  # every part of the feature workload scales with K
        t0 = time.perf_counter() - t_step
    per_row = max((time.perf_counter() - t0) / k, 1e-7)
    fixed = max(t_step - k * per_row, 0.0)
    est = max(64.0, (target_s - fixed) / per_row)
    return int(min(wl["K"], 1 << int(np.floor(np.log2(est)))))


def run_reference(args, wl):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    rng = np.random.default_rng(0)
    ks = cpu_sample_size(wl, 4.0)
    for _ in range(args.warmup):
        cpu_step(wl, ks, rng)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_step(wl, ks, rng)
    dt = (time.perf_counter() - t0) / args.steps
    value = ks / dt
    line = {
        "impl": "reference", "metric": "PPI update samples/sec", "value": value, "unit": "samples/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "config": {"workload": wl["desc"]},
        "cpu_baseline": {"value": value, "unit": "samples/s", "cores": 1, "kind": "port",
                         "host_cores": os.cpu_count(),
                         "sample": f"{ks} of {wl['K']} samples per step, full dimension; oracle port of the "
                                   "reference's NumPy path (its three-operand np.einsum is single-threaded; "
                                   "only the O(D^3) LAPACK tails use BLAS threads)"},
        "e2e": {"value": value, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    emit(line)


# ------------------------------------------------------------------------------------------------
# B200 arm
# ------------------------------------------------------------------------------------------------
def build_workload(wl, world, pinned=None):
    """-> (policy, solver, cost, K_total).  `pinned` = dict of pinned host arrays for the e2e leg."""
    from ppi_b200 import algorithms as alg, functions as fn, policies as pol, samplers as smp

    K_total = wl["K"] * world
    if wl["kind"] == "gaussian":
        D = wl["D"]
        mu = pinned["mu"] if pinned else np.ones(D)
        sigma = pinned["sigma"] if pinned else 0.5 * np.eye(D)
        policy = pol.GaussianPolicy(mu, sigma, smp.MonteCarlo(D, seed=1234))
        if wl.get("prefetch") is not None:
            policy.prefetch_normals = None if wl["prefetch"] == "none" else wl["prefetch"]
        solver = alg.Essps(n_elites=max(1, K_total // 64))
        cost = getattr(fn, wl["fn"])(D)
    elif wl["kind"] == "kernel":
        H, d = wl["H"], wl["d"]
        t = 0.01 * np.arange(H)
        mean, cin, cout = pol.design_moments(np.ones(d), -np.ones(d), ratio=1000.0)
        policy = pol.SquaredExponentialKernel(
            time_sequence=t, action_dimension=d, mean=mean, covariance_in=cin, covariance_out=cout,
            lengthscale=0.05, sampler=lambda dim: smp.MonteCarlo(dim, seed=1234),
            limiter=pol.Limiter(lower=-np.ones(d), upper=np.ones(d)))
        solver = alg.Lbps(delta=0.9)
        cost = fn.TrackingCost(t, d)
    else:
        T, d = wl["T"], wl["d"]
        t = np.linspace(0, 2, T)
        policy = pol.RffFeatures(
            time_sequence=t, action_dimension=d, mean=np.array([0.0, 1.5707]), covariance_in=np.array([1e2]),
            covariance_out=np.diag([1e-3, 1e-3]), lengthscale=np.sqrt(3e-2), order=10,
            sampler=lambda dim: smp.MonteCarlo(dim, seed=1234), use_derivatives=True, add_bias=True)
        policy.condition(np.zeros((1,)), np.array([[0.0, 1.5707]]))
        solver = alg.Essps(n_elites=10)
        cost = fn.TrackingCost(t, d)
    solver.reset(policy)
    return policy, solver, cost, K_total


def one_step(policy, solver, cost, K_total):
    actions, params = policy(K_total)
    costs = cost(actions)
    solver._stats = None
    solver.update(costs, params, policy)


def run_b200(args, wl):
    import torch
    import torch.distributed as dist

    from ppi_b200 import _lib

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py --impl b200 needs a GPU: the ppi_b200 path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    lib = _lib.load()
    if args.overlap is not None:
        _lib.call("ppi_set_option", _lib.OPT_GEMM_OVERLAP, args.overlap)
    if args.chunk is not None:
        _lib.call("ppi_set_option", _lib.OPT_GEMM_CHUNK, args.chunk)
    if args.pair is not None:
        _lib.call("ppi_set_option", _lib.OPT_GEMM_PAIR, args.pair)
    if args.prefetch_blocks is not None:
        _lib.call("ppi_set_option", _lib.OPT_PREFETCH_BLOCKS, args.prefetch_blocks)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    # ---- resident leg ----------------------------------------------------------------------------
    policy, solver, cost, K_total = build_workload(wl, world)
    step = lambda: one_step(policy, solver, cost, K_total)  # noqa: E731
    for _ in range(max(args.warmup, 1)):
        step()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches0 = lib.ppi_launch_count()
    lib.ppi_profile_enable(1)
    ms = timed(step, args.steps)
    prof = _lib.profile_read()
    lib.ppi_profile_enable(0)
    launches = lib.ppi_launch_count() - launches0
    clocks = sampler.stop() if rank == 0 else None
    ms_per_step = ms / args.steps
    value = K_total / (ms_per_step * 1e-3)

    # ---- end-to-end leg: host buffers in, host results out, every step ------------------------------
    h2d = d2h = 0
    if wl["kind"] == "gaussian":
        D = wl["D"]
        pin = {"mu": torch.ones(D, dtype=torch.float64).pin_memory(),
               "sigma": (0.5 * torch.eye(D, dtype=torch.float64)).pin_memory()}
        host = {k: v.numpy() for k, v in pin.items()}
        sink = {}

        def e2e_step():
            pol_, sol_, cost_, _ = build_workload(wl, world, pinned=host)
            one_step(pol_, sol_, cost_, K_total)
            sink["mu"], sink["sigma"] = pol_.mu, pol_.sigma  # device -> host read of the result

        h2d = D * D * 8 + D * 4
        d2h = D * D * 8 + 2 * D * 8 + _lib.STATS_BYTES
    else:
        sink = {}

        def e2e_step():
            pol_, sol_, cost_, _ = build_workload(wl, world)
            one_step(pol_, sol_, cost_, K_total)
            sink["mean"], sink["cov"] = pol_.mean, pol_.covariance_in

        n_in = policy.covariance_in.size + policy.mean.size
        h2d = 4 * (2 * n_in + policy.cov_out_sqrt.size)
        d2h = 8 * (policy.mean.size + 2 * policy.covariance_in.size) + _lib.STATS_BYTES
    for _ in range(2):
        e2e_step()
    e2e_ms = timed(e2e_step, args.steps) / args.steps
    e2e_value = K_total / (e2e_ms * 1e-3)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel ---------------------------------------------------------------
    pk = peaks()
    roofline = None
    if wl["kind"] == "gaussian" and "syrk_gemm" in prof:
        D = wl["D"]
        n, tot_ms = prof["syrk_gemm"]
        rows_per_launch = wl["K"] * args.steps / n        # samples contracted per launch
        algorithmic = 2.0 * rows_per_launch * D * D        # 2 D^2 useful flops per sample
        t = tot_ms / n * 1e-3
        scheme = lib.ppi_get_option(_lib.OPT_GEMM_SCHEME)
        # the policy's own Philox draws take the whitened path: A = nw z (fp16 pair) x B = z (exact fp16)
        white = scheme == _lib.GEMM_FP16X2
        products = 2 if white else 3
        name = ("syrk_white_kernel (tcgen05 kind::f16, whitened moments: in-kernel nw z as an fp16 pair x "
                "exact fp16 z, 2 partial products)" if white else "syrk_gemm_kernel<TF32X3> (tcgen05 kind::tf32)")
        roofline = {
            "kernel": name,
            "bound": "tensor", "achieved": algorithmic / t / 1e12, "peak": pk["bf16"], "unit": "TFLOP/s",
            "frac": algorithmic / t / 1e12 / pk["bf16"],
            "traffic": NCU_TRAFFIC.get(("syrk_white" if white else "syrk_pair", wl["D"], int(rows_per_launch))),
            "peak_source": f"dense 16-bit tensor peak, sustained, {pk['source']}",
            "algorithmic_flops_per_launch": algorithmic,
            "executed_tflops": executed_syrk_flops(rows_per_launch, D, products) / t / 1e12,
            "executed_note": f"MMA flops actually issued: {products} partial products on the upper-triangular "
                             "128x256 tiles (72 of 64 tile-equivalents at D = 2048); executed / peak is the "
                             "tensor-pipe utilisation",
            "launches": n, "avg_launch_ms": tot_ms / n,
            "share_of_step": tot_ms / (ms_per_step * args.steps),
        }
    else:
        # small-matrix workloads: every kernel but the (latency-bound) temperature solve streams the sample
        # tensor once; report the scope that takes the largest share of the step against the HBM roofline
        # with its compulsory bytes per sample
        if wl["kind"] == "kernel":
            x_bytes = 4.0 * wl["H"] * wl["d"]
            per_sample = {"sample_small": ("sample_matnorm_mma_kernel (write x)", x_bytes),
                          "cost": ("cost_tracking_kernel (read x)", x_bytes),
                          "wmoments_diag": ("wmoments_diag_kernel (read x)", x_bytes),
                          "wmoments_matnorm": ("wmoments_matnorm_mma_kernel (read x)", x_bytes)}
        else:
            F, dy = 21, 2 * wl["d"]
            y_bytes, w_bytes = 4.0 * wl["T"] * dy, 4.0 * F * wl["d"]
            per_sample = {"sample_small": ("feature_weights_kernel + features_expand_mma_kernel (write w twice, "
                                           "read it once, write y)", y_bytes + 3 * w_bytes),
                          "cost": ("cost_tracking_kernel (read y)", y_bytes),
                          "wmoments_diag": ("wmoments_diag_kernel (read w)", w_bytes),
                          "wmoments_matnorm": ("wmoments_matnorm_mma_kernel (read w)", w_bytes)}
        cand = [(prof[k][1], k) for k in per_sample if k in prof]
        if cand:
            _, key = max(cand)
            n, tot_ms = prof[key]
            name, nbytes = per_sample[key]
            t = tot_ms / n * 1e-3
            ach = nbytes * wl["K"] * args.steps / n / t / 1e9
            roofline = {"kernel": name, "bound": "hbm", "achieved": ach, "peak": pk["hbm"],
                        "unit": "GB/s", "frac": ach / pk["hbm"], "traffic": None, "peak_source": pk["source"],
                        "launches": n, "avg_launch_ms": tot_ms / n,
                        "share_of_step": tot_ms / (ms_per_step * args.steps),
                        "note": "fits L2 at the workload's own K; --samples 1048576 makes it HBM-resident"
                                if nbytes * wl["K"] < 126e6 else "sample tensor exceeds L2"}
    kernels = {k: {"launch_scopes": n, "ms_per_step": t / args.steps} for k, (n, t) in prof.items()}

    # ---- CPU baseline on a bounded sample ----------------------------------------------------------------
    cpu = None
    if world == 1 and not args.no_cpu:
        ks = cpu_sample_size(wl, 12.0)
        rng = np.random.default_rng(0)
        t0 = time.perf_counter()
        cpu_step(wl, ks, rng)
        dt = time.perf_counter() - t0
        cpu = {"value": ks / dt, "unit": "samples/s", "cores": 1, "kind": "port", "host_cores": os.cpu_count(),
               "sample": f"one update on {ks} of {wl['K']} samples at full dimension, {dt:.1f} s; oracle port of "
                         "the reference's NumPy path (naive three-operand einsum, single-threaded)"}
        if wl["kind"] == "gaussian":
            # the same math with BLAS contractions on every host core ("optimised CPU")
            kb = min(wl["K"], 16384)
            cpu_step(wl, 1024, rng, naive=False)
            t0 = time.perf_counter()
            cpu_step(wl, kb, rng, naive=False)
            dtb = time.perf_counter() - t0
            cpu["blas"] = {"value": kb / dtb, "unit": "samples/s", "cores": os.cpu_count(),
                           "sample": f"one update on {kb} samples, {dtb:.1f} s, BLAS restatement of the einsums "
                                     "(includes the K-independent O(D^3) factorisations)"}

    line = {
        "metric": "PPI update samples/sec", "value": value, "unit": "samples/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "dtype_note": "fp32 samples and results; dense products on tcgen05 as exact fp16 / fp16-pair partial "
                      "products with fp32 accumulation; fp64 scalars in the temperature solve and D x D tails",
        "data": "synthetic",
        "config": {"workload": wl["desc"], "samples_per_gpu": wl["K"], "samples_total": K_total,
                   "elements_per_s": value * (wl.get("D") or wl.get("H", 21) * wl["d"]),
                   "l2": "inputs exceed L2 (sample matrix per step >> 126 MB)" if wl["K"] * (wl.get("D") or 840) * 4 > 2 ** 28
                   else "working set fits L2; every step draws fresh Philox samples"},
        "e2e": {"value": e2e_value, "unit": "samples/s", "ms_per_step": e2e_ms, "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h},
        "gpu_launches": int(launches), "kernels": kernels, "roofline": roofline, "cpu_baseline": cpu,
        "clocks": clocks,
    }
    emit(line)
    if world > 1:
        dist.destroy_process_group()


def executed_syrk_flops(rows, D, products=3):
    """MMA flops actually issued: `products` partial products on the upper-triangular 128 x 256 tiles."""
    dp = (D + 127) // 128 * 128
    tiles_n = (dp + 255) // 256
    ntiles = sum(tiles_n - (tm * 128) // 256 for tm in range(dp // 128))
    return products * 2.0 * rows * ntiles * 128 * 256


def _claim_stdout():
    """Keep stdout for the one JSON line: libraries (NCCL prints its version banner to fd 1) are
    redirected to stderr; returns a file object on the original stdout."""
    sys.stdout.flush()
    real = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    return real


def emit(line):
    OUT.write(json.dumps(line) + "\n")
    OUT.flush()


OUT = sys.stdout


def main():
    global OUT
    OUT = _claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="C5", choices=sorted(WORKLOADS))
    ap.add_argument("--samples", type=int, default=None, help="samples per GPU (overrides the workload's K)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline leg")
    ap.add_argument("--chunk", type=int, default=None, help="override PPI_OPT_GEMM_CHUNK")
    ap.add_argument("--pair", type=int, default=None, choices=[0, 1], help="override PPI_OPT_GEMM_PAIR")
    ap.add_argument("--overlap", type=int, default=None, choices=[0, 1],
                    help="override PPI_OPT_GEMM_OVERLAP (operand staging on a side stream)")
    ap.add_argument("--prefetch", default=None, choices=["none", "sample", "update"],
                    help="override GaussianPolicy.prefetch_normals (base normals of the next draw generated "
                         "ahead on a side stream)")
    ap.add_argument("--prefetch-blocks", type=int, default=None, help="override PPI_OPT_PREFETCH_BLOCKS")
    args = ap.parse_args()
    wl = dict(WORKLOADS[args.workload])
    wl["prefetch"] = args.prefetch
    if args.samples:
        wl["K"] = args.samples
        wl["desc"] += f" [K overridden to {args.samples}]"
    if args.impl == "reference":
        run_reference(args, wl)
    else:
        run_b200(args, wl)


if __name__ == "__main__":
    main()
