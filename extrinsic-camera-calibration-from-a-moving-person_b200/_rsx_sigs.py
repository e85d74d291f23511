"""ctypes signatures of include/rsx.h (declared on the library object loaded by _lib.lib())."""
import ctypes as C

vp, i32, i64, f64 = C.c_void_p, C.c_int, C.c_int64, C.c_double
P_f64, P_i64, P_i32, P_u64 = (C.POINTER(C.c_double), C.POINTER(C.c_int64), C.POINTER(C.c_int32),
                              C.POINTER(C.c_uint64))

SIGS = {
    "rsx_create": (i32, [i32, C.POINTER(vp)]),
    "rsx_destroy": (i32, [vp]),
    "rsx_sample_stream": (i32, [P_u64, i64, i64, i64, P_i32, P_i32]),
    "rsx_problem_set": (i32, [vp, i32, i64, i64, vp, vp, vp]),
    "rsx_fit_score": (i32, [vp, vp, vp, i64, vp, vp, vp, vp]),
    "rsx_prune": (i32, [vp, i32, vp]),
    "rsx_prune_stats": (i32, [vp, P_i64, P_f64, vp]),
    "rsx_argmin": (i32, [vp, vp, i64, P_i64, P_f64, vp]),
    "rsx_argmin_static": (i32, [i32, vp, i64, vp, vp]),
    "rsx_winner_masks": (i32, [vp, vp, vp, f64, f64, vp, vp, vp, vp, P_i64, vp]),
    "rsx_calib_linear": (i32, [vp, vp, vp, vp, vp, P_i64, vp]),
    "rsx_calib_linear_eigs": (i32, [vp, P_f64]),
    "rsx_triangulate": (i32, [vp, vp, vp, vp]),
    "rsx_fit_stats": (i32, [vp, P_i64, vp]),
    "rsx_timing_enable": (i32, [vp, i32]),
    "rsx_timing_read": (i32, [vp, P_f64, P_i64]),
}


def declare(L):
    for name, (res, args) in SIGS.items():
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args


def exported_symbols():
    return list(SIGS.keys())
