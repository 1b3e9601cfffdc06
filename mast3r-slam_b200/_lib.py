"""ctypes binding of the C ABI declared in include/m3s.h (libm3s_b200.so, built in-tree by _build.py).

There is no CPU or library fallback: if the shared library is missing, or no CUDA device is
visible when a compute entry point is called, a RuntimeError is raised.
"""
from __future__ import annotations

import ctypes as C
import os

_PKG = os.path.dirname(os.path.abspath(__file__))
# M3S_LIB: development override to A/B-test an alternative build of the same C ABI
SO_PATH = os.environ.get("M3S_LIB") or os.path.join(_PKG, "libm3s_b200.so")

GN_RAYS, GN_POINTS, GN_CALIB = 0, 1, 2
GN_EDGE_STRIDE = 36
GN_PREC_DEFAULT, GN_PREC_F64, GN_PREC_MIXED = 0, 1, 2


class GNParams(C.Structure):
    """struct m3s_gn_params (include/m3s.h)."""
    _fields_ = [
        ("mode", C.c_int32), ("max_iter", C.c_int32), ("pin", C.c_int32),
        ("pixel_border", C.c_int32), ("img_w", C.c_int32), ("img_h", C.c_int32),
        ("sigma_a", C.c_double), ("sigma_b", C.c_double), ("C_thresh", C.c_double),
        ("Q_thresh", C.c_double), ("delta_thresh", C.c_double), ("z_eps", C.c_double),
        ("K4", C.c_float * 4), ("precision", C.c_int32), ("semantics", C.c_int32),
    ]


TRACK_RAY_DIST, TRACK_CALIB = 0, 1


class TrackParams(C.Structure):
    """struct m3s_track_params (include/m3s.h)."""
    _fields_ = [
        ("mode", C.c_int32), ("max_iters", C.c_int32), ("img_w", C.c_int32), ("img_h", C.c_int32),
        ("pixel_border", C.c_int32), ("reserved", C.c_int32),
        ("sigma_ray", C.c_double), ("sigma_dist", C.c_double), ("sigma_pixel", C.c_double),
        ("sigma_depth", C.c_double), ("huber", C.c_double), ("rel_error", C.c_double),
        ("delta_norm", C.c_double), ("depth_eps", C.c_double), ("K4", C.c_float * 4),
    ]


_vp, _i, _f, _sz, _i64 = C.c_void_p, C.c_int, C.c_float, C.c_size_t, C.c_int64

# name -> (restype, argtypes); every symbol include/m3s.h declares
SIGNATURES = {
    "m3s_version": (_i, []),
    "m3s_last_error": (C.c_char_p, []),
    "m3s_device_count": (_i, []),
    "m3s_sm_count": (_i, [_i]),
    "m3s_prep_iter_proj": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _vp]),
    "m3s_iter_proj_workspace_bytes": (_sz, [_i]),
    "m3s_iter_proj": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _f, _f, _vp, _sz, _vp]),
    "m3s_refine_matches": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "m3s_refine_matches_f32": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "m3s_match_tail": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _f, _vp]),
    "m3s_iter_proj_metal": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _f, _f, _vp]),
    "m3s_refine_matches_metal": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "m3s_set_path_counters": (_i, [_vp]),
    "m3s_match_simple": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _f, _vp]),
    "m3s_match_workspace_bytes": (_sz, [_i, _i, _i, _i]),
    "m3s_match_iterative_proj": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _f,
                                      _f, _i, _i, _f, _vp, _sz, _vp]),
    "m3s_match_iterative_proj_static": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i,
                                             _f, _f, _i, _i, _f, _vp]),
    "m3s_match_iterative_proj_host": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i, _i, _f,
                                           _f, _i, _i, _f, _i, _i]),
    "m3s_match_host_release": (_i, [_i]),
    "m3s_constrain_points_to_ray": (_i, [_vp, C.POINTER(C.c_float), _vp, _i, _i, _i, _vp]),
    "m3s_edge_confidence": (_i, [_vp, _vp, _vp, _vp, _f, _vp, _vp, _vp, _i, C.c_longlong, _vp]),
    "m3s_track_gate": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _f, _f, _vp, _vp, _vp, _vp, _i, _vp]),
    "m3s_track_workspace_bytes": (_sz, [_i]),
    "m3s_track_opt_pose": (_i, [C.POINTER(TrackParams), _vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _vp, _vp,
                                _vp, _vp, _sz, _vp]),
    "m3s_gn_plan_create": (_vp, [_vp, _vp, _i, _i, _i64, _i, _i]),
    "m3s_gn_plan_destroy": (None, [_vp]),
    "m3s_gn_plan_num_free": (_i, [_vp]),
    "m3s_gn_plan_dim": (_i, [_vp]),
    "m3s_gn_plan_ld": (_i, [_vp]),
    "m3s_gn_plan_factor_path": (_i, [_vp]),
    "m3s_gn_plan_sparse_info": (_i, [_vp, _vp]),
    "m3s_gn_plan_edge_blocks": (_vp, [_vp, C.POINTER(_i64)]),
    "m3s_gn_plan_H": (_vp, [_vp]),
    "m3s_gn_plan_g": (_vp, [_vp]),
    "m3s_gn_plan_dx": (_vp, [_vp]),
    "m3s_gn_plan_poses": (_vp, [_vp]),
    "m3s_gn_plan_timeline": (_vp, [_vp, C.POINTER(C.c_int64)]),
    "m3s_gn_set_poses": (_i, [_vp, _vp, _vp]),
    "m3s_gn_get_poses": (_i, [_vp, _vp, _vp]),
    "m3s_gn_accumulate": (_i, [_vp, C.POINTER(GNParams), _vp, _vp, _vp, _vp, _vp, _i, _i, _vp]),
    "m3s_gn_assemble": (_i, [_vp, _vp]),
    "m3s_gn_solve_update": (_i, [_vp, C.POINTER(GNParams), _vp]),
    "m3s_gn_exchange_create": (_i, [_vp, _i, _i, _vp]),
    "m3s_gn_exchange_attach": (_i, [_vp, _vp]),
    "m3s_gn_exchange_detach": (_i, [_vp]),
    "m3s_gn_exchange_release": (_i, [_vp]),
    "m3s_gn_status": (_i, [_vp, C.POINTER(C.c_int32), C.POINTER(C.c_double), _vp]),
    "m3s_gauss_newton": (_i, [_vp, C.POINTER(GNParams), _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
}

_lib = None


class M3SError(RuntimeError):
    pass


def load():
    """Load libm3s_b200.so and declare every prototype.  Raises if the library is not built."""
    global _lib
    if _lib is None:
        if not os.path.exists(SO_PATH):
            raise M3SError(
                f"{SO_PATH} not found: build it with `python -c 'import __graft_entry__ as g; "
                "g.build()'` (nvcc, sm_100a).  There is no CPU fallback.")
        lib = C.CDLL(SO_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = load().m3s_last_error().decode(errors="replace")
        raise M3SError(f"{what or 'm3s'} failed (rc={rc}): {msg}")


def require_device() -> int:
    """Number of visible CUDA devices; raises (loudly, no fallback) when there is none."""
    n = load().m3s_device_count()
    if n <= 0:
        raise M3SError("no CUDA device visible: the m3s-b200 kernels have no CPU fallback")
    return n
