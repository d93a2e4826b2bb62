#!/usr/bin/env python
"""bench.py -- PPI update samples/sec on B200 (BASELINE.json metric), one JSON line on stdout.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
                  [--workload C5|C2|C3|C4] [--scaling strong|weak] [--normals fp16|fp32]

A "step" is one pass of the hot path (ppi/algorithms.py:85-88): sample K actions from the prior ->
cost -> Gibbs-posterior temperature + weights -> weighted moment match -> new prior.  Default
workload = config C5 of BASELINE.json: GaussianPolicy, D = 2048, K = 2^20 samples IN TOTAL, split
evenly over the N ranks (strong scaling, SURVEY 8(d) C5 / 8(e): rank r owns rows [r K/N, (r+1) K/N),
costs are all-gathered, whitened moments all-reduced), Rastrigin cost, ESSPS temperature with
N = K/64, on-device Philox normals.  --scaling weak keeps K = 2^20 PER GPU (round 1's mode).

value      whole-job samples/s with the policy state resident in HBM (CUDA events, max over ranks)
e2e        same step through the public API with HOST buffers: a fresh policy is built from pinned
           host mu / sigma every step and the fitted mu / sigma are read back to the host
roofline   the dominant kernel (tcgen05 covariance GEMM): algorithmic flops per launch / its
           CUDA-event time measured inside the timed region, against the measured 16-bit tensor peak
cpu_baseline  the oracle port (NumPy restatement of the reference, its naive einsum) on a bounded
           sample of the same workload, timed on this box's host cores (rank 0, N = 1 only)
extra      compact records of the side workloads C2 / C3 / C4 (and C3 / C4 with K = 2^20, where the
           sample tensor lives in HBM) with their own clock samples (N = 1 only)
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
    # K = samples of the whole job
    "C5": dict(kind="gaussian", D=2048, K=2 ** 20, fn="Rastrigin",
               desc="C5: GaussianPolicy D=2048, K=2^20 samples, Rastrigin, Essps(N=K/64), Philox"),
    "C2": dict(kind="gaussian", D=1024, K=65536, fn="Rastrigin",
               desc="C2: GaussianPolicy D=1024, K=65536, Rastrigin, Essps(N=K/64), Philox"),
    "C3": dict(kind="kernel", H=30, d=28, K=4096,
               desc="C3: SquaredExponentialKernel H=30 d=28 (door-v0 shapes), K=4096, tracking return, Lbps(0.9)"),
    "C4": dict(kind="features", T=1000, d=2, K=65536,
               desc="C4: RffFeatures T=1000 F=21 d=2 (ball-in-a-cup shapes), K=65536, tracking return, Essps(10)"),
}


# dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, from the committed
# `ncu --set full` capture under profiles/ (see profiles/README.md); keyed by (kernel, D, samples per launch).
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
        return self

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def mark(self):
        """Number of samples seen so far (to cut the record at the start of a timed region)."""
        return len(self.lines)

    def stop(self, first=0):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"], "samples": 0}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        return self.summary(first)

    def summary(self, first=0, last=None):
        sm, smax, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.lines[first:last]:
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
def host_threads():
    """Cores this process may run on (the cgroup / affinity mask, not the host's core count), at most 32."""
    try:
        n = len(os.sched_getaffinity(0))
    except AttributeError:
        n = os.cpu_count() or 1
    return max(1, min(32, n))


def set_blas_threads(n=None):
    """Pin the BLAS / OpenMP pools explicitly (torchrun exports OMP_NUM_THREADS=1 to its workers, which
    silently changed the O(D^3) LAPACK share of the CPU arm between N = 1 and N > 1 in round 1)."""
    n = host_threads() if n is None else int(n)
    try:
        from threadpoolctl import threadpool_limits

        threadpool_limits(limits=n)
        return n
    except Exception:
        return None


def log(msg):
    sys.stderr.write(f"[bench {time.strftime('%H:%M:%S')}] {msg}\n")
    sys.stderr.flush()


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


def cpu_fit(wl, sizes, rng, naive=True, repeats=1):
    """Time one update at each sample size and fit t(K) = fixed + per_row K (the CPU arm
    of a workload whose full K needs hours is a linear extrapolation in K).  -> dict with the per-size
    timings, the fitted coefficients and the samples/s the fit predicts at the workload's own K."""
    times = []
    for k in sizes:
        best = None
        for _ in range(repeats):
            t0 = time.perf_counter()
            cpu_step(wl, k, rng, naive=naive)
            dt = time.perf_counter() - t0
            best = dt if best is None else min(best, dt)
        times.append(best)
    ks = np.asarray(sizes, np.float64)
    ts = np.asarray(times, np.float64)
    if len(sizes) > 1:
        per_row, fixed = np.polyfit(ks, ts, 1)
        per_row, fixed = max(float(per_row), 1e-12), max(float(fixed), 0.0)
    else:
        per_row, fixed = float(ts[0] / ks[0]), 0.0
    k_full = float(wl["K"])
    return {"sizes": [int(k) for k in sizes], "seconds": [float(t) for t in times], "per_row_s": per_row,
            "fixed_s": fixed, "value_at_full_K": k_full / (fixed + per_row * k_full)}


def cpu_sizes(wl, budget_s):
    """Three sample sizes (k, 2k, 4k) whose naive updates together take about budget_s seconds.  The per-row
    cost is timed on the K-proportional part alone (the moment-matching / sampling einsums); the
    K-independent part (O(D^3) factorisations, ~1 s at D = 2048 and noisy) is what is left of a whole step."""
    from oracle import ppi_oracle as O

    rng = np.random.default_rng(0)
    cpu_step(wl, 8, rng)  # warm caches / lazy imports
    k = 16
    t0 = time.perf_counter()
    if wl["kind"] == "gaussian":
        O.m_projection(np.zeros(k), rng.standard_normal((k, wl["D"])))
    elif wl["kind"] == "kernel":
        x = rng.standard_normal((k, wl["H"], wl["d"]))
        O.sample_matnorm(x, np.zeros(wl["d"]), np.zeros((wl["H"], wl["d"])), np.eye(wl["H"]), np.eye(wl["d"]))
        O.m_projection_mavn(np.zeros(k), x, np.eye(wl["H"]), np.eye(wl["d"]))
    else:
        cpu_step(wl, k, rng)      # every part of the feature workload scales with K
    per_row = max((time.perf_counter() - t0) / k, 1e-6)
    t0 = time.perf_counter()
    cpu_step(wl, k, rng)
    fixed = max(time.perf_counter() - t0 - k * per_row, 0.0)
    base = (budget_s - 3 * fixed) / (7 * per_row)          # k + 2k + 4k rows
    base = int(max(16, min(wl["K"] // 4, 1 << int(np.floor(np.log2(max(base, 16)))))))
    # hard bound: never more than budget_s of per-row work, whatever the timings above said
    while base > 16 and 7 * base * per_row > 1.5 * budget_s:
        base //= 2
    return [base, 2 * base, 4 * base]


def config_of(wl, K_total, world, scaling):
    """The `config` object -- identical in the b200 and the reference arm."""
    return {"workload": wl["desc"], "samples_total": int(K_total), "scaling": scaling,
            "sharding": f"K split evenly over {world} rank(s)" if scaling == "strong"
            else f"K = {wl['K']} per rank x {world} rank(s)"}


def run_reference(args, wl):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    world = max(1, args.gpus)
    K_total = wl["K"] * (world if args.scaling == "weak" else 1)
    threads = set_blas_threads()
    # every step = one bounded update at one of three sample sizes (k, 2k, 4k, cycling); the value is the
    # samples/s of the full-K update predicted by the linear fit t(K) = fixed + per_row K
    sizes = cpu_sizes(wl, max(6.0, 150.0 / max(1, args.steps + args.warmup) * 3))
    rng = np.random.default_rng(0)
    for i in range(args.warmup):
        cpu_step(wl, sizes[i % 3], rng)
    per_size = {k: [] for k in sizes}
    t_all = time.perf_counter()
    for i in range(args.steps):
        k = sizes[i % 3]
        t0 = time.perf_counter()
        cpu_step(wl, k, rng)
        per_size[k].append(time.perf_counter() - t0)
    wall = time.perf_counter() - t_all
    ks = [k for k in sizes if per_size[k]]
    ts = [float(np.mean(per_size[k])) for k in ks]
    if len(ks) > 1:
        per_row, fixed = np.polyfit(np.asarray(ks, float), np.asarray(ts), 1)
        per_row, fixed = max(float(per_row), 1e-12), max(float(fixed), 0.0)
    else:
        per_row, fixed = ts[0] / ks[0], 0.0
    value = K_total / (fixed + per_row * K_total)
    blas = None
    if wl["kind"] == "gaussian":
        kb = 4096
        cpu_step(wl, 512, rng, naive=False)
        fit = cpu_fit(wl, [kb, 2 * kb], rng, naive=False)
        blas = {"value": K_total / (fit["fixed_s"] + fit["per_row_s"] * K_total), "unit": "samples/s",
                "cores": threads, "fit": fit,
                "sample": "BLAS restatement of the einsums ((nw x)^T x, z L^T) on every host core, linear fit in K"}
    line = {
        "impl": "reference", "metric": "PPI update samples/sec", "value": value, "unit": "samples/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": wall / max(1, args.steps) * 1e3,
        "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f64",
        "data": "synthetic", "config": config_of(wl, K_total, world, args.scaling),
        "cpu_baseline": {"value": value, "unit": "samples/s", "cores": 1, "kind": "port",
                         "host_cores": os.cpu_count(), "blas_threads": threads,
                         "fit": {"sizes": ks, "mean_seconds": ts, "per_row_s": per_row, "fixed_s": fixed},
                         "blas": blas,
                         "sample": f"steps cycle over updates on {sizes} of {K_total} samples at full dimension; "
                                   "value = K / (fixed + per_row K) from the linear fit over those sizes; oracle "
                                   "port of the reference's NumPy path (its three-operand np.einsum is "
                                   "single-threaded; only the O(D^3) LAPACK tails use the BLAS threads)"},
        "e2e": {"value": value, "unit": "samples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    emit(line)


# ------------------------------------------------------------------------------------------------
# B200 arm
# ------------------------------------------------------------------------------------------------
def build_workload(wl, K_total, pinned=None):
    """-> (policy, solver, cost).  `pinned` = dict of pinned host arrays for the e2e leg."""
    from ppi_b200 import algorithms as alg, functions as fn, policies as pol, samplers as smp

    if wl["kind"] == "gaussian":
        D = wl["D"]
        mu = pinned["mu"] if pinned else np.ones(D)
        sigma = pinned["sigma"] if pinned else 0.5 * np.eye(D)
        sampler = smp.MonteCarlo(D, seed=1234)
        if wl.get("normals") == "fp32":
            sampler = smp.MonteCarlo(D, seed=1234, full_precision=True)
        policy = pol.GaussianPolicy(mu, sigma, sampler)
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
    return policy, solver, cost


def one_step(policy, solver, cost, K_total):
    actions, params = policy(K_total)
    costs = cost(actions)
    solver._stats = None
    solver.update(costs, params, policy)


class Harness:
    """torch / distributed plumbing shared by the main measurement and the extras."""

    def __init__(self):
        import torch
        import torch.distributed as dist

        self.torch, self.dist = torch, dist
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py --impl b200 needs a GPU: the ppi_b200 path has no CPU fallback")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        if self.world > 1:
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            dist.init_process_group("nccl", device_id=self.dev)

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def timed(self, fn, steps):
        torch = self.torch
        self.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        self.barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=self.dev, dtype=torch.float64)
        if self.world > 1:
            self.dist.all_reduce(ms, op=self.dist.ReduceOp.MAX)
        return float(ms.item())

    def close(self):
        if self.world > 1:
            self.dist.destroy_process_group()


def measure(hx, wl, K_total, steps, warmup, clock=None):
    """Resident-state leg of one workload: -> dict(ms_per_step, value, launches, kernels, scopes, clocks)."""
    from ppi_b200 import _lib, device

    lib = _lib.load()
    policy, solver, cost = build_workload(wl, K_total)
    step = lambda: one_step(policy, solver, cost, K_total)  # noqa: E731
    for _ in range(max(warmup, 3)):
        step()
    first = clock.mark() if clock is not None else 0
    launches0 = lib.ppi_launch_count()
    lib.ppi_profile_enable(1)
    device.profile_scopes(True)
    ms = hx.timed(step, steps)
    prof = _lib.profile_read()
    scopes = device.profile_scopes_read()
    lib.ppi_profile_enable(0)
    device.profile_scopes(False)
    launches = lib.ppi_launch_count() - launches0
    clocks = clock.summary(first) if clock is not None else None
    ms_per_step = ms / steps
    kernels = {k: {"launch_scopes": n, "ms_per_step": t / steps} for k, (n, t) in prof.items()}
    kernels.update({k: {"launch_scopes": n, "ms_per_step": t / steps, "host_scope": True}
                    for k, (n, t) in scopes.items()})
    return dict(ms_per_step=ms_per_step, value=K_total / (ms_per_step * 1e-3), launches=int(launches), prof=prof,
                kernels=kernels, clocks=clocks, policy=policy,
                nominal_path=int(getattr(policy, "n_regularized", 0)) == 0)


def measure_fresh(hx, wl, K_total, budget_s, chunk, clock):
    """Side workloads: `chunk` timed steps at a time on a FRESH policy (hundreds of consecutive updates would
    collapse the search distribution and time the revert branch), repeated until budget_s of device time
    have been sampled.  -> same dict as measure()."""
    from ppi_b200 import _lib, device

    lib = _lib.load()
    first = clock.mark() if clock is not None else 0
    total_ms, steps, launches, nominal = 0.0, 0, 0, True
    prof_acc, scope_acc = {}, {}
    t_wall = time.perf_counter()
    while total_ms < budget_s * 1e3 and steps < 20000 and time.perf_counter() - t_wall < 20.0 * budget_s + 10.0:
        policy, solver, cost = build_workload(wl, K_total)
        step = lambda: one_step(policy, solver, cost, K_total)  # noqa: E731
        step()
        launches0 = lib.ppi_launch_count()
        lib.ppi_profile_enable(1)
        device.profile_scopes(True)
        total_ms += hx.timed(step, chunk)
        for acc, src in ((prof_acc, _lib.profile_read()), (scope_acc, device.profile_scopes_read())):
            for k, (n, t) in src.items():
                a = acc.setdefault(k, [0, 0.0])
                a[0] += n
                a[1] += t
        lib.ppi_profile_enable(0)
        device.profile_scopes(False)
        launches += lib.ppi_launch_count() - launches0
        steps += chunk
        nominal = nominal and int(getattr(policy, "n_regularized", 0)) == 0
    clocks = clock.summary(first) if clock is not None else None
    prof = {k: (n, t) for k, (n, t) in prof_acc.items()}
    kernels = {k: {"launch_scopes": n, "ms_per_step": t / steps} for k, (n, t) in prof.items()}
    kernels.update({k: {"launch_scopes": n, "ms_per_step": t / steps, "host_scope": True}
                    for k, (n, t) in scope_acc.items()})
    ms_per_step = total_ms / steps
    return dict(ms_per_step=ms_per_step, value=K_total / (ms_per_step * 1e-3), launches=int(launches), prof=prof,
                kernels=kernels, clocks=clocks, steps=steps, nominal_path=nominal)


def small_roofline(wl, K_local, prof, steps, ms_per_step, pk):
    """Small-matrix workloads: every kernel but the (latency-bound) temperature solve streams the sample
    tensor once; report the scope that takes the largest share of the step against the HBM roofline with
    its compulsory bytes per sample."""
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
    if not cand:
        return None
    _, key = max(cand)
    n, tot_ms = prof[key]
    name, nbytes = per_sample[key]
    t = tot_ms / n * 1e-3
    ach = nbytes * K_local * steps / n / t / 1e9
    hbm = {k: round(per_sample[k][1] * K_local * steps / prof[k][0] / (prof[k][1] / prof[k][0] * 1e-3) / 1e9 / pk["hbm"], 3)
           for k in per_sample if k in prof}
    return {"kernel": name, "bound": "hbm", "achieved": ach, "peak": pk["hbm"],
            "unit": "GB/s", "frac": ach / pk["hbm"], "traffic": None, "peak_source": pk["source"],
            "launches": n, "avg_launch_ms": tot_ms / n,
            "share_of_step": tot_ms / (ms_per_step * steps), "hbm_frac_by_scope": hbm,
            "note": "fits L2 at the workload's own K; K = 2^20 makes it HBM-resident"
                    if nbytes * K_local < 126e6 else "sample tensor exceeds L2"}


def run_b200(args, wl):
    from ppi_b200 import _lib

    hx = Harness()
    torch, world, rank = hx.torch, hx.world, hx.rank
    lib = _lib.load()
    for key, val in ((_lib.OPT_GEMM_OVERLAP, args.overlap), (_lib.OPT_GEMM_CHUNK, args.chunk),
                     (_lib.OPT_GEMM_PAIR, args.pair), (_lib.OPT_PREFETCH_BLOCKS, args.prefetch_blocks),
                     (_lib.OPT_SAMPLE_CHUNK, args.sample_chunk)):
        if val is not None:
            _lib.call("ppi_set_option", key, val)

    K_total = wl["K"] * (world if args.scaling == "weak" else 1)
    K_local = K_total // world
    clock = ClockSampler(hx.local).start() if rank == 0 else None

    # ---- resident leg ----------------------------------------------------------------------------
    log(f"resident leg: {wl['desc']}, K_total = {K_total}, world = {world}")
    res = measure(hx, wl, K_total, args.steps, args.warmup, clock)
    log(f"resident leg done: {res['ms_per_step']:.3f} ms/step")
    ms_per_step, value, prof = res["ms_per_step"], res["value"], res["prof"]

    # ---- end-to-end leg: host buffers in, host results out, every step ------------------------------
    policy = res["policy"]
    if wl["kind"] == "gaussian":
        D = wl["D"]
        # a DENSE, well-conditioned host covariance (spectrum in [0.5, 1.5]): a diagonal one would hand the
        # tensor cores operands full of zeros, which run cooler and faster on a power-capped board than the
        # dense factors the resident leg sees after its warm-up
        gen = torch.Generator().manual_seed(20261105)
        a = torch.randn((D, D), dtype=torch.float64, generator=gen)
        pin = {"mu": torch.ones(D, dtype=torch.float64).pin_memory(),
               "sigma": (0.5 * torch.eye(D, dtype=torch.float64) + (a @ a.T) * (0.25 / D)).pin_memory()}
        del a
        host = {k: v.numpy() for k, v in pin.items()}
        sink = {}

        def e2e_step():
            pol_, sol_, cost_ = build_workload(wl, K_total, pinned=host)
            one_step(pol_, sol_, cost_, K_total)
            sink["mu"], sink["sigma"] = pol_.mu, pol_.sigma  # device -> host read of the result

        h2d = D * D * 8 + D * 4
        d2h = D * D * 8 + 2 * D * 8 + _lib.STATS_BYTES
    else:
        sink = {}

        def e2e_step():
            pol_, sol_, cost_ = build_workload(wl, K_total)
            one_step(pol_, sol_, cost_, K_total)
            sink["mean"], sink["cov"] = pol_.mean, pol_.covariance_in

        n_in = policy.covariance_in.size + policy.mean.size
        h2d = 4 * (2 * n_in + policy.cov_out_sqrt.size)
        d2h = 8 * (policy.mean.size + 2 * policy.covariance_in.size) + _lib.STATS_BYTES
    del policy
    res.pop("policy")
    for _ in range(2):
        e2e_step()
    e2e_ms = hx.timed(e2e_step, args.steps) / args.steps
    log(f"e2e leg done: {e2e_ms:.3f} ms/step")
    e2e_value = K_total / (e2e_ms * 1e-3)

    if rank != 0:
        hx.close()
        return

    # ---- roofline of the dominant kernel ---------------------------------------------------------------
    pk = peaks()
    roofline = None
    if wl["kind"] == "gaussian" and "syrk_gemm" in prof:
        D = wl["D"]
        n, tot_ms = prof["syrk_gemm"]
        rows_per_launch = K_local * args.steps / n        # samples contracted per launch
        algorithmic = 2.0 * rows_per_launch * D * D        # 2 D^2 useful flops per sample
        t = tot_ms / n * 1e-3
        scheme = lib.ppi_get_option(_lib.OPT_GEMM_SCHEME)
        # the policy's own Philox draws take the whitened path: A = nw z (fp16 pair) x B = z (exact fp16)
        white = scheme == _lib.GEMM_FP16X2 and wl.get("normals") != "fp32"
        products = 2 if white else 3
        name = ("syrk_white_kernel (tcgen05 kind::f16, whitened moments: in-kernel nw z as an fp16 pair x "
                "exact fp16 z, 2 partial products)" if white else "syrk_gemm_kernel (tcgen05, 3 partial products)")
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
        roofline = small_roofline(wl, K_local, prof, args.steps, ms_per_step, pk)

    # ---- side workloads (N = 1): compact records with their own clock samples --------------------------
    extra = None
    if world == 1 and not args.no_extras and args.workload == "C5":
        extra = {}
        for name, base, K in (("C2", "C2", None), ("C3", "C3", None), ("C4", "C4", None),
                              ("C3_K1M", "C3", 1 << 20), ("C4_K1M", "C4", 1 << 20)):
            w2 = dict(WORKLOADS[base])
            if K:
                w2["K"] = K
            try:
                log(f"extra {name}")
                r2 = measure_fresh(hx, w2, w2["K"], 0.7, 4, clock)       # >= 0.7 s of device time: >= 20 clock samples
                rl = (small_roofline(w2, w2["K"], r2["prof"], r2["steps"], r2["ms_per_step"], pk)
                      if w2["kind"] != "gaussian" else None)
                extra[name] = {"workload": w2["desc"] + (f" [K = {K}]" if K else ""), "steps": r2["steps"],
                               "ms_per_step": r2["ms_per_step"], "value": r2["value"], "unit": "samples/s",
                               "kernels_ms_per_step": {k: round(v["ms_per_step"], 5) for k, v in r2["kernels"].items()},
                               "roofline": rl, "clocks": r2["clocks"], "gpu_launches": r2["launches"],
                               "nominal_path": r2["nominal_path"]}
            except Exception as e:  # a side workload must never cost the headline line
                extra[name] = {"error": repr(e)[:300]}
            torch.cuda.empty_cache()
    clocks = clock.stop() if clock is not None else None
    if clocks is not None:
        clocks = dict(res["clocks"], all_samples=clocks["samples"])

    # ---- CPU baseline on a bounded sample ----------------------------------------------------------------
    cpu = None
    if world == 1 and not args.no_cpu:
        threads = set_blas_threads()
        log(f"cpu baseline: {threads} BLAS threads")
        sizes = cpu_sizes(wl, 14.0)
        log(f"cpu baseline sizes {sizes}")
        fit = cpu_fit(wl, sizes, np.random.default_rng(0))
        cpu = {"value": fit["value_at_full_K"], "unit": "samples/s", "cores": 1, "kind": "port",
               "host_cores": os.cpu_count(), "blas_threads": threads, "fit": fit,
               "sample": f"one update each on {sizes} of {wl['K']} samples at full dimension "
                         f"({sum(fit['seconds']):.1f} s); value = K / (fixed + per_row K) from the linear fit; oracle "
                         "port of the reference's NumPy path (naive three-operand einsum, single-threaded)"}
        if wl["kind"] == "gaussian":
            # the same math with BLAS contractions on every host core ("optimised CPU")
            rng = np.random.default_rng(0)
            log("cpu baseline: BLAS restatement")
            cpu_step(wl, 512, rng, naive=False)
            fb = cpu_fit(wl, [2048, 4096], rng, naive=False)
            cpu["blas"] = {"value": fb["value_at_full_K"], "unit": "samples/s", "cores": threads, "fit": fb,
                           "sample": "BLAS restatement of the einsums on every host core, linear fit in K"}

    line = {
        "metric": "PPI update samples/sec", "value": value, "unit": "samples/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
        "scaling": args.scaling, "vs_baseline": None, "dtype": "f32",
        "dtype_note": "fp32 samples and results; dense products on tcgen05 as exact fp16 / fp16-pair partial "
                      "products with fp32 accumulation; fp64 scalars in the temperature solve and D x D tails",
        "data": "synthetic",
        "config": config_of(wl, K_total, world, args.scaling),
        "details": {"samples_per_gpu": K_local, "normals": wl.get("normals") or "fp16-grid Philox (default)",
                    "elements_per_s": value * (wl.get("D") or wl.get("H", 21) * wl["d"]),
                    "nominal_path": res["nominal_path"],
                    "l2": "inputs exceed L2 (sample matrix per step >> 126 MB)"
                    if K_local * (wl.get("D") or 840) * 4 > 2 ** 28
                    else "working set fits L2; every step draws fresh Philox samples"},
        "e2e": {"value": e2e_value, "unit": "samples/s", "ms_per_step": e2e_ms, "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": d2h},
        "gpu_launches": res["launches"], "kernels": res["kernels"], "roofline": roofline, "cpu_baseline": cpu,
        "clocks": clocks, "extra": extra,
    }
    emit(line)
    hx.close()


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
    ap.add_argument("--scaling", default="strong", choices=["strong", "weak"],
                    help="strong (default): the workload's K is split over the ranks; weak: K per rank")
    ap.add_argument("--normals", default=None, choices=["fp16", "fp32"],
                    help="fp32: unquantised Philox normals through the injected-normal (pair-store) path -- what "
                         "the step costs without rounding the base normals to the fp16 grid")
    ap.add_argument("--samples", type=int, default=None, help="total samples (overrides the workload's K)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline leg")
    ap.add_argument("--no-extras", action="store_true", help="skip the C2 / C3 / C4 side records")
    ap.add_argument("--chunk", type=int, default=None, help="override PPI_OPT_GEMM_CHUNK")
    ap.add_argument("--sample-chunk", type=int, default=None, help="override PPI_OPT_SAMPLE_CHUNK")
    ap.add_argument("--pair", type=int, default=None, choices=[0, 1], help="override PPI_OPT_GEMM_PAIR")
    ap.add_argument("--overlap", type=int, default=None, choices=[0, 1],
                    help="override PPI_OPT_GEMM_OVERLAP (operand staging on a side stream)")
    ap.add_argument("--prefetch", default=None, choices=["none", "sample", "moments", "update"],
                    help="override GaussianPolicy.prefetch_normals (base normals of the next draw generated "
                         "ahead on a side stream)")
    ap.add_argument("--prefetch-blocks", type=int, default=None, help="override PPI_OPT_PREFETCH_BLOCKS")
    args = ap.parse_args()
    wl = dict(WORKLOADS[args.workload])
    wl["prefetch"] = args.prefetch
    wl["normals"] = args.normals if args.normals == "fp32" else None
    if args.samples:
        wl["K"] = args.samples
        wl["desc"] += f" [K overridden to {args.samples}]"
    if args.impl == "reference":
        run_reference(args, wl)
    else:
        run_b200(args, wl)


if __name__ == "__main__":
    main()
