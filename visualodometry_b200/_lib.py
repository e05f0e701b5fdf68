"""ctypes binding of the C ABI in include/b2ba.h (libb2ba.so).  No CPU fallback: if the CUDA
library is missing or no GPU is present, using the engine raises."""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libb2ba.so")

OK, E_INVALID, E_CUDA, E_NCCL, E_DUPLICATE, E_STATE, E_UNSUPPORTED = 0, -1, -2, -3, -4, -5, -6
MAX_ITERS, CONVERGED, NUMERICAL_FAILURE = 1, 2, 3
MEM_HOST, MEM_DEVICE = 0, 1
NOALLOC_MAX_POINTS = 32768     # include/b2ba.h B2BA_NOALLOC_MAX_POINTS
D = dict(PM_PERM=1, PT_PTR=2, CM_PERM=3, CAM_PTR=4, FIDX=5, BLK_PTR=6, BLK_COL=7, SCHED=8,
         RESIDUALS=16, WEIGHTS=17, JC=18, JP=19,
         B=32, GC=33, C=34, GP=35, S=36, GRED=37, DC=38, DP=39, SCALARS=40,
         HIST_COST=64, HIST_MU=65, HIST_ACCEPT=66, POSE_HIST_COST=80, POSE_HIST_MU=81, POSE_HIST_ACCEPT=82)
# every symbol include/b2ba.h declares
SYMBOLS = ["b2ba_default_options", "b2ba_create", "b2ba_destroy", "b2ba_last_error", "b2ba_set_problem",
           "b2ba_set_problem_batch", "b2ba_set_state", "b2ba_reset", "b2ba_solve", "b2ba_get_summaries", "b2ba_get_timings",
           "b2ba_get", "b2ba_optimize_batch", "b2ba_refine_poses", "b2ba_track_pose", "b2ba_ransac_pose", "b2ba_triangulate", "b2ba_reproj_errors", "b2ba_cull", "b2ba_dump", "b2ba_create_comm", "b2ba_nccl_unique_id", "b2ba_set_point_range",
           "b2ba_peak_flops", "b2ba_solve_window_noalloc", "b2ba_version"]


class Options(C.Structure):
    _fields_ = [("max_iters", C.c_int32), ("loss", C.c_int32), ("loss_scale", C.c_double),
                ("z_min", C.c_double), ("init_mu", C.c_double), ("ftol", C.c_double),
                ("gtol", C.c_double), ("ptol", C.c_double), ("solver", C.c_int32),
                ("pcg_max_iters", C.c_int32), ("pcg_tol", C.c_double), ("flags", C.c_int32),
                ("pad_", C.c_int32)]


class Summary(C.Structure):
    _fields_ = [("status", C.c_int32), ("iterations", C.c_int32), ("accepted", C.c_int32),
                ("rejected", C.c_int32), ("n_invalid", C.c_int32), ("pcg_iters", C.c_int32),
                ("initial_cost", C.c_double), ("final_cost", C.c_double), ("final_mu", C.c_double)]


class Timings(C.Structure):
    _fields_ = [("setup_ms", C.c_double), ("solve_ms", C.c_double), ("phase_ms", C.c_double * 8),
                ("iterations_enqueued", C.c_int32), ("pad_", C.c_int32)]


PHASES = ("linearise", "schur", "exchange", "solve", "backsub", "control")

_lib = None


def load():
    """Load libb2ba.so; raises (never falls back) if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} not built: run `python -m visualodometry_b200._build` "
                           "(the b2ba engine is CUDA-only; there is no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    vp, i32, i64, dp = C.c_void_p, C.c_int32, C.c_int64, C.POINTER(C.c_double)
    lib.b2ba_default_options.argtypes = [C.POINTER(Options)]
    lib.b2ba_default_options.restype = None
    lib.b2ba_create.argtypes = [C.POINTER(vp), C.c_int, C.POINTER(Options)]
    lib.b2ba_destroy.argtypes = [vp]
    lib.b2ba_destroy.restype = None
    lib.b2ba_last_error.argtypes = [vp]
    lib.b2ba_last_error.restype = C.c_char_p
    lib.b2ba_set_problem.argtypes = [vp, i32, i32, i64, vp, vp, vp, vp, vp, vp, dp, C.c_int]
    lib.b2ba_set_problem_batch.argtypes = [vp, i32, vp, vp, vp, vp, vp, vp, vp, vp, vp, dp, C.c_int]
    lib.b2ba_set_state.argtypes = [vp, vp, vp, C.c_int]
    lib.b2ba_reset.argtypes = [vp]
    lib.b2ba_solve.argtypes = [vp, i32, vp, vp]
    lib.b2ba_get_summaries.argtypes = [vp, vp]
    lib.b2ba_get_timings.argtypes = [vp, C.POINTER(Timings)]
    lib.b2ba_get.argtypes = [vp, vp, vp, C.c_int]
    lib.b2ba_optimize_batch.argtypes = [vp, i32, vp, vp, vp, vp, vp, vp, vp, vp, vp, dp, i32, vp, vp, vp, i32]
    lib.b2ba_refine_poses.argtypes = [vp, i32, vp, vp, vp, vp, dp, i32, vp, i32]
    lib.b2ba_track_pose.argtypes = [vp, i64, vp, vp, vp, dp, C.c_double, i32, i32, vp, vp, vp, i32]
    lib.b2ba_ransac_pose.argtypes = [vp, i64, vp, vp, vp, dp, C.c_double, i32, i32, C.c_uint64, i32, vp, vp, vp, vp, i32]
    lib.b2ba_triangulate.argtypes = [vp, i32, dp, dp, vp, vp, C.c_double, C.c_double, vp, vp, i32]
    lib.b2ba_reproj_errors.argtypes = [vp, vp, i32]
    lib.b2ba_cull.argtypes = [vp, C.c_double, vp, vp, vp, C.POINTER(i64), i32, i32]
    lib.b2ba_dump.argtypes = [vp, C.c_int, vp, i64]
    lib.b2ba_dump.restype = i64
    lib.b2ba_create_comm.argtypes = [vp, vp, C.c_int, C.c_int]
    lib.b2ba_nccl_unique_id.argtypes = [vp]
    lib.b2ba_set_point_range.argtypes = [vp, i32, i32]
    lib.b2ba_peak_flops.argtypes = [C.c_int, dp, dp]
    lib.b2ba_solve_window_noalloc.argtypes = [vp, vp, vp, vp, i32, vp]
    lib.b2ba_version.restype = C.c_int
    _lib = lib
    return lib
