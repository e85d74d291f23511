"""ctypes binding of libbax.so (include/bax.h, include/rsx.h).  There is NO CPU fallback: if the
library is missing or a call fails, an exception is raised."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libbax.so")

_lib = None

vp = C.c_void_p
i32 = C.c_int
i64 = C.c_int64
f64 = C.c_double
P_f64 = C.POINTER(C.c_double)
P_i64 = C.POINTER(C.c_int64)
P_i32 = C.POINTER(C.c_int32)

TRF_MAX_HIST = 64


class TrfState(C.Structure):
    """bax_trf_state (include/bax.h)."""
    _fields_ = [("ftol", f64), ("xtol", f64), ("gtol", f64), ("max_nfev", i64),
                ("x", vp), ("x_new", vp), ("step", vp),
                ("cost", f64), ("g_inf", f64), ("g_2", f64), ("x_norm", f64), ("h_scale", f64), ("Delta", f64),
                ("alpha", f64), ("initial_cost", f64),
                ("nfev", i64), ("njev", i64), ("iteration", i64), ("n_trials", i64),
                ("status", C.c_int32), ("n_hist", C.c_int32),
                ("hist", (f64 * 8) * TRF_MAX_HIST)]


class BatchArgs(C.Structure):
    """bax_batch_args (include/bax.h)."""
    PP = C.POINTER(C.c_void_p)
    _fields_ = [("n_problems", C.c_int32), ("n_workers", C.c_int32), ("device", C.c_int32),
                ("C", C.c_int32), ("N", C.c_int32), ("J", C.c_int32), ("J3", C.c_int32), ("B", C.c_int32),
                ("K", PP), ("y2d", PP), ("s2d", PP), ("p3d", PP), ("s3d", PP), ("bones", P_i32),
                ("lambda1", f64), ("lambda2", f64), ("R0", PP), ("t0", PP), ("rvec0", PP),
                ("ftol", f64), ("xtol", f64), ("gtol", f64), ("max_nfev", i64),
                ("theta_out", vp), ("info_out", vp), ("blocking_sync", C.c_int32)]


_SIGS = {
    "bax_last_error": (C.c_char_p, []),
    "bax_version": (i32, []),
    "bax_create": (i32, [i32, C.POINTER(vp)]),
    "bax_destroy": (i32, [vp]),
    "bax_set_blocking_sync": (i32, [vp, i32]),
    "bax_problem_set": (i32, [vp, i32, i32, i32, i32, i32, vp, vp, vp, vp, vp, P_i32, f64, f64, vp]),
    "bax_problem_dims": (i32, [vp, P_i64]),
    "bax_observations_rebind": (i32, [vp, vp, vp]),
    "bax_residual": (i32, [vp, vp, vp, vp]),
    "bax_residual_jacobian": (i32, [vp, vp, vp, vp, vp, vp, vp, vp]),
    "bax_prior_jacobian": (i32, [vp, vp, vp, vp, vp]),
    "bax_reduce_buffer": (i32, [vp, i32, C.POINTER(vp), P_i64]),
    "bax_set_global_frames": (i32, [vp, i64]),
    "bax_cost_begin": (i32, [vp, vp, vp]),
    "bax_cost_mid": (i32, [vp, vp]),
    "bax_cost_finish": (i32, [vp, P_f64, vp]),
    "bax_linearize_begin": (i32, [vp, vp, vp]),
    "bax_linearize_mid": (i32, [vp, vp]),
    "bax_linearize_post": (i32, [vp, vp]),
    "bax_linearize_finish": (i32, [vp, P_f64, vp]),
    "bax_linearize_push": (i32, [vp]),
    "bax_linearize_pop": (i32, [vp]),
    "bax_trial_begin": (i32, [vp, f64, vp]),
    "bax_trial_mid": (i32, [vp, vp, i32, vp]),
    "bax_trial_finish": (i32, [vp, P_f64, vp]),
    "bax_trial_phi_prime": (i32, [vp, vp]),
    "bax_trf_begin": (i32, [vp, C.POINTER(TrfState), vp]),
    "bax_trf_iterate": (i32, [vp, C.POINTER(TrfState), vp]),
    "bax_trf_solve": (i32, [vp, C.POINTER(TrfState), vp]),
    "bax_batch_solve": (i32, [C.POINTER(BatchArgs)]),
    "bax_batch_release": (i32, []),
    "bax_comm_unique_id": (i32, [vp]),
    "bax_comm_create": (i32, [i32, vp, i32, i32, C.POINTER(vp)]),
    "bax_comm_destroy": (i32, [vp]),
    "bax_allreduce_attach": (i32, [vp, vp]),
    "bax_comm_stats": (i32, [vp, P_i64]),
    "bax_triangulate": (i32, [vp, vp, vp, vp, P_i32, vp]),
    "bax_reproject_error": (i32, [vp, vp, vp, vp, vp, vp]),
    "bax_timing_enable": (i32, [vp, i32]),
    "bax_timing_read": (i32, [vp, P_f64, P_i64]),
    "bax_timing_reset": (i32, [vp]),
}

BUF_BONE, BUF_COST, BUF_LIN, BUF_SCHUR, BUF_RHS2, BUF_STEP, BUF_NORM, BUF_GMAX, BUF_RHS2_HEAD = range(9)
T_NAMES = ["rj", "build", "syrk", "dense", "backsub", "linearize", "cost", "other"]


class BaxError(RuntimeError):
    pass


def lib():
    """Load libbax.so (once).  Raises if it has not been built — there is no fallback path."""
    global _lib
    if _lib is None:
        if not os.path.isfile(LIB_PATH):
            raise BaxError(f"{LIB_PATH} is missing: build it with __graft_entry__.build() "
                           "(csrc/build.sh); this package has no CPU fallback")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGS.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        try:
            from . import _rsx_sigs
            _rsx_sigs.declare(L)
        except ImportError:
            pass
        _lib = L
    return _lib


def check(rc, what=""):
    if rc != 0:
        msg = lib().bax_last_error()
        raise BaxError(f"{what} failed ({rc}): {msg.decode() if msg else ''}")


def exported_symbols():
    return list(_SIGS.keys())
