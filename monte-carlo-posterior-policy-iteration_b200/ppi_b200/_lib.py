"""ctypes binding of the C ABI in include/ppi_b200.h (libppi_b200.so, built by ../build.py).

There is no fallback: if the shared library is missing or a call fails, an exception is raised.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libppi_b200.so")

# enums of include/ppi_b200.h
KERNEL_KINDS = {"se": 0, "matern12": 1, "matern32": 2, "matern52": 3, "periodic": 4, "white": 5}
FN_IDS = {"Himmelblau": 0, "Rosenbrock": 1, "Styblinski": 2, "Rastrigin": 3, "NoisySphere": 4}
OPT_GEMM_SCHEME, OPT_GEMM_CHUNK, OPT_GEMM_OVERLAP, OPT_GEMM_PAIR, OPT_PREFETCH_BLOCKS = 0, 1, 2, 3, 4
OPT_SAMPLE_CHUNK = 5
GEMM_TF32X3, GEMM_FP16X2 = 0, 1
TEMP_FIXED_NORMALISED, TEMP_FIXED_SHIFTED, TEMP_REPS_KL, TEMP_ESS_TARGET, TEMP_LOWER_BOUND = range(5)


class WeightStats(C.Structure):
    """struct ppi_weight_stats"""

    _fields_ = [
        ("alpha", C.c_double), ("lse", C.c_double), ("ess", C.c_double),
        ("weighted_cost", C.c_double), ("weight_entropy", C.c_double),
        ("cost_min", C.c_double), ("cost_max", C.c_double), ("cost_mean", C.c_double),
        ("cost_std", C.c_double), ("objective", C.c_double),
        ("n_valid", C.c_int64), ("argmax", C.c_int64),
        ("all_nan", C.c_int32), ("iterations", C.c_int32), ("status", C.c_int32), ("pad", C.c_int32),
    ]


STATS_BYTES = C.sizeof(WeightStats)

_p, _i, _i64, _u64, _d, _sz = C.c_void_p, C.c_int, C.c_int64, C.c_uint64, C.c_double, C.c_size_t

# name -> (restype, argtypes); mirrors include/ppi_b200.h one to one
SIGNATURES = {
    "ppi_version": (_i, []),
    "ppi_last_error": (C.c_char_p, []),
    "ppi_set_option": (_i, [_i, _i]),
    "ppi_get_option": (_i, [_i]),
    "ppi_philox_normal": (_i, [_p, _i64, _i64, _u64, _u64, _u64, _p]),
    "ppi_philox_normal_full": (_i, [_p, _i64, _i64, _u64, _u64, _u64, _p]),
    "ppi_kernel_gram_chol": (_i, [_p, _i, _p, _i, _i, _d, _d, _d, _p, _p, _p, _p, _p, _p]),
    "ppi_cholesky_small": (_i, [_p, _i, _p, _p, _p, _p]),
    "ppi_sobol_normal": (_i, [_p, _i64, _i64, _p, _p, _i, _u64, _d, _p]),
    "ppi_cubature_points": (_i, [_p, _i, _p]),
    "ppi_matnorm_update_workspace_bytes": (_sz, [_i, _i]),
    "ppi_matnorm_update": (_i, [_p, _p, _i64, _i, _i, _p, _p, _d, _i64, _p, _p, _p, _p, _p, _sz, _p]),
    "ppi_kernel_shift_workspace_bytes": (_sz, [_i, _i]),
    "ppi_kernel_shift": (_i, [_i, _d, _d, _d, _p, _i, _p, _i, _p, _p, _p, _p, _i, _p, _p, _p, _d, _d, _p, _p,
                              _p, _p, _p, _p]),
    "ppi_sample_matnorm": (_i, [_p, _u64, _u64, _u64, _i64, _i, _i, _p, _p, _p, _p, _i, _p, _p, _p, _p]),
    "ppi_sample_features": (_i, [_p, _u64, _u64, _u64, _i64, _i, _i, _i, _p, _p, _p, _p, _p, _p, _p, _p,
                                 _p, _p, _p]),
    "ppi_sample_iid": (_i, [_p, _u64, _u64, _u64, _i64, _i, _i, _p, _p, _p, _p, _p, _p, _p]),
    "ppi_sample_gaussian_workspace_bytes": (_sz, [_i64, _i64]),
    "ppi_sample_gaussian": (_i, [_p, _u64, _u64, _u64, _i64, _i64, _p, _p, _p, _p, _sz, _p]),
    "ppi_zstore_bytes": (_sz, [_i64, _i64, _i]),
    "ppi_philox_prefetch": (_i, [_u64, _u64, _u64, _i64, _i64, _p, _sz, _p]),
    "ppi_sample_gaussian_pre": (_i, [_i64, _i64, _p, _p, _p, _p, _p, _sz, _p]),
    "ppi_philox_normal_cq": (_i, [_p, _i64, _i64, _u64, _u64, _u64, _p]),
    "ppi_sample_gaussian_keep": (_i, [_p, _u64, _u64, _u64, _i64, _i64, _p, _p, _p, _p, _sz, _p, _sz, _p]),
    "ppi_sample_gaussian_cost_workspace_bytes": (_sz, [_i64, _i64]),
    "ppi_sample_gaussian_cost": (_i, [_p, _u64, _u64, _u64, _i64, _i64, _p, _p, _p, _p, _sz, _i, _i, _p, _p, _sz, _p]),
    "ppi_wmoments_white_workspace_bytes": (_sz, [_i64, _i64, _i]),
    "ppi_wmoments_white": (_i, [_p, _i, _p, _i64, _i64, _p, _p, _p, _i, _p, _sz, _p]),
    "ppi_cost_function": (_i, [_i, _p, _i64, _i64, _p, _p, _p, _p]),
    "ppi_cost_tracking": (_i, [_p, _i64, _i, _i, _p, _i, _p, _p]),
    "ppi_solve_temperature_workspace_bytes": (_sz, [_i64]),
    "ppi_solve_temperature": (_i, [_p, _i64, _i, _d, _p, _p, _p, _p, _sz, _p]),
    "ppi_solve_temperature_static": (_i, [_p, _i64, _i, _d, _p, _p, _p, _p]),
    "ppi_normalise_log_weights": (_i, [_p, _i64, _p, _p, _p, _sz, _p]),
    "ppi_cem_workspace_bytes": (_sz, [_i64]),
    "ppi_cem_weights": (_i, [_p, _i64, _i64, _p, _p, _p, _p, _sz, _p]),
    "ppi_wmoments_diag": (_i, [_p, _p, _i64, _i64, _p, _p, _p, _p, _p, _p]),
    "ppi_wmoments_matnorm": (_i, [_p, _p, _i64, _i, _i, _p, _p, _p, _p]),
    "ppi_wmoments_dense_workspace_bytes": (_sz, [_i64, _i64]),
    "ppi_wmoments_dense": (_i, [_p, _p, _i64, _i64, _p, _p, _p, _i, _p, _sz, _p]),
    "ppi_finalize_gaussian": (_i, [_p, _p, _i64, _p, _i, _p, _p, _p]),
    "ppi_finalize_matnorm": (_i, [_p, _p, _i, _i, _d, _p, _p, _p, _p, _p, _p, _p, _p]),
    "ppi_dense_workspace_bytes": (_sz, [_i64]),
    "ppi_cholesky_blocked": (_i, [_p, _i64, _p, _p, _p, _p, _sz, _p]),
    "ppi_white_finish": (_i, [_p, _p, _p, _p, _p, _i64, _p, _p, _p, _p, _p, _sz, _p]),
    "ppi_tri_gram": (_i, [_p, _i64, _p, _p, _p]),
    "ppi_tri_matvec": (_i, [_p, _i64, _p, _p, _p, _p]),
    "ppi_tri_solve_sumsq": (_i, [_p, _i64, _p, _p, _p, _p, _sz, _p]),
    "ppi_launch_count": (_i64, []),
    "ppi_profile_enable": (_i, [_i]),
    "ppi_profile_read": (_i, [_i, C.POINTER(_i64), C.POINTER(_d)]),
}

PROFILE_SLOTS = {"syrk_gemm": 0, "sample_gemm": 1, "syrk_operand": 2, "sample_operand": 3,
                 "wmoments_diag": 4, "cost": 5, "solve": 6, "syrk_reduce": 7, "sample_small": 8,
                 "wmoments_matnorm": 9, "dense_tail": 10}


def profile_read():
    """{slot name: (scopes, total device ms)} of the CUDA-event brackets since ppi_profile_enable(1)."""
    out = {}
    for name, slot in PROFILE_SLOTS.items():
        n, ms = _i64(0), _d(0.0)
        call("ppi_profile_read", slot, C.byref(n), C.byref(ms))
        if n.value:
            out[name] = (n.value, ms.value)
    return out

_lib = None


class PpiError(RuntimeError):
    pass


def load():
    """Load libppi_b200.so (once).  Raises if it has not been built: there is no CPU fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise PpiError(
            f"{LIB_PATH} not found: build it with `python monte-carlo-posterior-policy-iteration_b200/"
            "build.py` (or __graft_entry__.build()).  ppi_b200 has no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def call(name, *args):
    """Call an int-returning entry point and raise PpiError on a non-zero status."""
    lib = load()
    rc = getattr(lib, name)(*args)
    if rc != 0:
        msg = lib.ppi_last_error().decode(errors="replace")
        raise PpiError(f"{name} failed with status {rc}: {msg}")


def query(name, *args):
    return getattr(load(), name)(*args)
