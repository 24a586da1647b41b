"""ctypes binding of libb200ba.so (C ABI in include/b200ba.h).

There is no CPU fallback: if the shared library is missing, or no CUDA device is
visible when an engine is created, this raises.  The library is built in-tree by
``__graft_entry__.build()`` (``make -C sfm-python_b200/csrc``).
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# B200BA_LIB: load another build of the same library (kernel A/B experiments, tools/ab_variants.sh); default in-tree build
LIB_PATH = os.environ.get("B200BA_LIB") or os.path.join(_HERE, "libb200ba.so")

STATUS = {0: "OK", -1: "BAD_ARG", -2: "CUDA", -3: "NCCL", -4: "NOT_FINITE", -5: "OOM", -6: "INDEX_RANGE",
          -7: "NO_PROBLEM"}
SOLVERS = {"auto": 0, "pcg": 1, "cholesky": 2, "chol": 2}


class Options(C.Structure):
    _fields_ = [("ftol", C.c_double), ("xtol", C.c_double), ("gtol", C.c_double), ("max_nfev", C.c_int64),
                ("verbose", C.c_int32), ("solver", C.c_int32), ("lambda0", C.c_double), ("pcg_rtol", C.c_double),
                ("pcg_max_iter", C.c_int32), ("fixed_iters", C.c_int32)]


class Result(C.Structure):
    _fields_ = [("cost", C.c_double), ("initial_cost", C.c_double), ("optimality", C.c_double), ("rms", C.c_double),
                ("lambda_final", C.c_double), ("nfev", C.c_int64), ("njev", C.c_int64), ("status", C.c_int32),
                ("n_iter", C.c_int32), ("pcg_iters", C.c_int64), ("solve_ms", C.c_double)]


class Stats(C.Structure):
    _fields_ = [(n, C.c_double) for n in ("ms_cam_prep", "ms_linearize_points", "ms_camera_blocks", "ms_schur",
                                          "ms_allreduce", "ms_camera_solve", "ms_backsub_trial", "ms_residual")] + \
               [(n, C.c_int64) for n in ("launches", "n_linearize", "n_pcg_iters", "schur_entries", "schur_blocks",
                                         "bytes_device")] + \
               [(n, C.c_double) for n in ("ms_lin_fused", "ms_slot_reduce")] + \
               [(n, C.c_int64) for n in ("n_tiles", "n_pair_slots", "n_self_slots")] + [("ms_jacobian", C.c_double),
                                                                                         ("n_tl_builds", C.c_int64),
                                                                                         ("ms_reproj", C.c_double),
                                                                                         ("ms_residual_kernel", C.c_double),
                                                                                         ("ms_chol_mc", C.c_double)]


_P = C.c_void_p
_D = C.POINTER(C.c_double)
_I64 = C.POINTER(C.c_int64)

# name -> (restype, argtypes); every symbol declared in include/b200ba.h
SIGNATURES = {
    "b200ba_version": (C.c_int, []),
    "b200ba_create": (C.c_int, [C.POINTER(_P), C.c_int]),
    "b200ba_create_multi": (C.c_int, [C.POINTER(_P), C.POINTER(C.c_int), C.c_int]),
    "b200ba_destroy": (C.c_int, [_P]),
    "b200ba_last_error": (C.c_char_p, [_P]),
    "b200ba_nccl_unique_id": (C.c_int, [_P]),
    "b200ba_comm_init": (C.c_int, [_P, _P, C.c_int, C.c_int]),
    "b200ba_set_problem": (C.c_int, [_P, C.c_int64, C.c_int64, C.c_int64, _P, _P, _P, _P]),
    "b200ba_residuals": (C.c_int, [_P, _P, _P, _D]),
    "b200ba_residuals_dev": (C.c_int, [_P, _P, _P, _D]),
    "b200ba_residuals_noalloc": (C.c_int, [_P, C.c_int64, C.c_int64, C.c_int64, _P, _P, _P, _P, _P, _P, _P]),
    "b200ba_pack_cameras": (C.c_int, [_P, C.c_int64, _P, _P]),
    "b200ba_unpack_cameras": (C.c_int, [_P, C.c_int64, _P, _P]),
    "b200ba_solve_pnp": (C.c_int, [_P, C.c_int64, _P, _P, _P, _P, _P, C.c_int32, _P]),
    "b200ba_triangulate": (C.c_int, [_P, C.c_int64, _P, _P, _P, _P, _P, _P, _P, _P]),
    "b200ba_triangulate_nonlinear": (C.c_int, [_P, C.c_int64, _P, _P, _P, _P, C.c_int32, _P]),
    "b200ba_reproj_errors": (C.c_int, [_P, _P, _P, _P, _D]),
    "b200ba_reproj_errors_dev": (C.c_int, [_P, _P, _P, _P, _D]),
    "b200ba_jacobian_blocks": (C.c_int, [_P, _P, _P, _P]),
    "b200ba_jacobian_blocks_dev": (C.c_int, [_P, _P, _P, _P, _P]),
    "b200ba_normal_equations": (C.c_int, [_P, _P, C.c_double, _P, _P, _P, _D]),
    "b200ba_linearize_dev": (C.c_int, [_P, _P, C.c_double]),
    "b200ba_lm_step": (C.c_int, [_P, _P, C.c_double, C.c_int, _P, _D, _D]),
    "b200ba_solve": (C.c_int, [_P, _P, C.POINTER(Options), C.POINTER(Result)]),
    "b200ba_host_alloc": (_P, [C.c_int64]),
    "b200ba_host_free": (None, [_P]),
    "b200ba_solve_io": (C.c_int, [_P, _P, _P, C.POINTER(Options), C.POINTER(Result)]),
    "b200ba_solve_dev": (C.c_int, [_P, _P, C.POINTER(Options), C.POINTER(Result)]),
    "b200ba_set_two_level": (C.c_int, [_P, C.c_int]),
    "b200ba_set_loss": (C.c_int, [_P, C.c_int, C.c_double]),
    "b200ba_set_fixed_cameras": (C.c_int, [_P, _P, C.c_int64]),
    "b200ba_get_stats": (C.c_int, [_P, C.POINTER(Stats)]),
    "b200ba_reset_stats": (C.c_int, [_P]),
    "b200ba_set_profiling": (C.c_int, [_P, C.c_int]),
    "b200ba_fp64_peak": (C.c_int, [_P, _D]),
    "b200ba_stream": (_P, [_P]),
    "b200ba_debug_fetch": (C.c_int64, [_P, C.c_char_p, _P, C.c_int64]),
}

_lib = None


def load():
    """Load libb200ba.so and bind every entry point; raises if the library has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(there is no CPU fallback)")
    lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib
