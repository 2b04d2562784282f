"""ctypes binding of libptgs.so (include/ptgs.h).  The product path has no CPU
fallback: if the library is missing and cannot be built, importing fails."""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libptgs.so")

PTGS_F32, PTGS_F64 = 0, 1
PTGS_OK = 0
PTGS_E_CAPACITY = -3
PTGS_E_RANGE = -4
CAMERA_FLOATS = 20
EXTREME_DIRS = 8                   # PTGS_EXTREME_DIRS (checked against the library at load)
POLY_FLOATS = 4 + 3 * EXTREME_DIRS
PAIR_RECORD_BYTES = 24
COMPACT_XY_STATIC_MAX_POINTS = 1 << 26    # PTGS_COMPACT_XY_STATIC_MAX_POINTS

_vp, _i32, _i64, _f64, _sz = C.c_void_p, C.c_int32, C.c_int64, C.c_double, C.c_size_t

# name -> (restype, argtypes); mirrors include/ptgs.h one to one
SIGNATURES = {
    "ptgs_version": (C.c_int, []),
    "ptgs_extreme_dirs": (C.c_int, []),
    "ptgs_last_error": (C.c_char_p, []),
    "ptgs_launch_count": (_i64, []),
    "ptgs_reset_launch_count": (None, []),
    "ptgs_aos_to_soa_f32": (C.c_int, [_vp, _i64, _vp, _vp, _vp, _vp]),
    "ptgs_aos_to_soa_f64": (C.c_int, [_vp, _i64, _vp, _vp, _vp, _vp]),
    "ptgs_manhattan_apply": (C.c_int, [_vp, _vp, _vp, _i64, C.POINTER(_f64), _vp, _vp, _vp, _vp]),
    "ptgs_aos_to_soa_manhattan": (C.c_int, [_vp, _i64, C.POINTER(_f64), _vp, _vp, _vp, _vp]),
    "ptgs_radius_outlier_workspace_bytes": (_sz, [_i64]),
    "ptgs_radius_outlier_mask": (C.c_int, [C.c_int, _vp, _vp, _vp, _vp, _i64, _f64, C.c_int, _vp, _vp,
                                           _vp, _sz, _vp]),
    "ptgs_slab_select_workspace_bytes": (_sz, [_i64]),
    "ptgs_slab_select": (C.c_int, [C.c_int, _vp, _vp, _vp, _i64, _f64, _f64, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "ptgs_slab_keep": (C.c_int, [C.c_int, _vp, _vp, _vp, _i64, _f64, _f64, _vp, _vp]),
    "ptgs_compact_xy_workspace_bytes": (_sz, [_i64]),
    "ptgs_compact_xy": (C.c_int, [C.c_int, _vp, _vp, _vp, _i64, _vp, _vp, _vp, _vp, _sz, _vp]),
    "ptgs_compact_xy_static": (C.c_int, [C.c_int, _vp, _vp, _vp, _i64, _vp, _vp, _vp, _vp]),
    "ptgs_bounds_xy": (C.c_int, [C.c_int, _vp, _vp, _vp, _i64, _vp, _vp]),
    "ptgs_balanced_split_workspace_bytes": (_sz, [C.c_int, _i64, _f64, C.c_int]),
    "ptgs_balanced_split": (C.c_int, [C.c_int, _vp, _vp, _vp, _i64, _vp, _f64, C.c_int, _f64,
                                      _vp, _vp, _vp, _vp, _vp, _i32, _vp, _vp, _sz, _vp]),
    "ptgs_expand_workspace_bytes": (_sz, [_i64, _i32]),
    "ptgs_expand_count": (C.c_int, [C.c_int, _vp, _vp, _i64, _vp, _i32, _vp, _vp, _sz, _vp]),
    "ptgs_expand_fill": (C.c_int, [C.c_int, _vp, _vp, _i64, _vp, _i32, _vp, _vp, _vp, _i64, _vp,
                                   _sz, _vp]),
    "ptgs_expand_count_grid": (C.c_int, [C.c_int, _vp, _vp, _i64, _vp, _i32, C.POINTER(C.c_double), _vp, _vp, _vp,
                                         _vp, _sz, _vp]),
    "ptgs_expand_fill_grid": (C.c_int, [C.c_int, _vp, _vp, _i64, _vp, _i32, C.POINTER(C.c_double), _vp, _vp, _vp,
                                        _vp, _vp, _i64, _vp, _sz, _vp]),
    "ptgs_assign_cameras": (C.c_int, [_vp, _i32, _vp, _i32, _vp, _vp, _i64, _vp]),
    "ptgs_voxel_downsample_workspace_bytes": (_sz, [_i64, _i32]),
    "ptgs_voxel_downsample": (C.c_int, [C.c_int, _vp, _vp, _vp, _vp, _vp, _vp, _i64, _i32, _f64, _vp,
                                        _vp, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "ptgs_visibility_workspace_bytes": (_sz, [_i32]),
    "ptgs_visibility_init": (C.c_int, [_i64, _vp, _vp, _vp, _vp, _vp]),
    "ptgs_visibility_count": (C.c_int, [_vp, _vp, _vp, _vp, _i32, _vp, _vp, _vp, _i32, _vp, _vp,
                                        _vp, _vp, _vp, _sz, _vp]),
    "ptgs_visibility_sample": (C.c_int, [_vp, _vp, _vp, _vp, _i32, _i32, _vp, _vp, _vp, _vp, _i64, _vp]),
    "ptgs_visibility_sample_count": (C.c_int, [_vp, _vp, _vp, _vp, _i32, _vp, _vp, _vp, _i64, _vp, _vp, _vp, _vp]),
    "ptgs_visibility_krect": (C.c_int, [_vp, _vp, _vp, _vp, _i32, _vp, _vp, _vp, _vp, _vp, _i64, _f64,
                                        _vp, _vp]),
    "ptgs_visibility_bounds": (C.c_int, [_vp, _vp, _vp, _vp, _i32, _vp, _vp, _vp, _vp, _vp, _vp,
                                         _i64, _i32, _f64, _f64, _i32, _vp, _vp, _vp, _vp, _vp, _vp,
                                         _vp, _vp, _vp]),
    "ptgs_visibility_plan": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, _i64, _i32, _i64, _vp, _vp]),
    "ptgs_visibility_candidates": (C.c_int, [_vp, _vp, _vp, _vp, _i32, _vp, _vp, _vp, _i32, _vp, _vp,
                                             _vp, _vp, _vp, _vp, _sz, _vp]),
    "ptgs_visibility_hull_workspace_bytes": (_sz, [_i64]),
    "ptgs_visibility_hull": (C.c_int, [_vp, _vp, _vp, _vp, _i64, _vp, _vp, _i32, _f64, _f64, _vp,
                                       _vp, _vp, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "ptgs_select_cameras": (C.c_int, [_vp, _vp, _vp, _i32, _vp, _vp, _vp, _vp]),
    "ptgs_points3d_index": (_i64, [_vp, _i64, _vp, _i64]),
    "ptgs_points3d_decode": (C.c_int, [_vp, _vp, _i64, _vp, _vp, _vp, _vp, _vp, _vp]),
    "ptgs_pack_records": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _i64, _vp, _vp]),
    "ptgs_unpack_records": (C.c_int, [_vp, _i32, _i64, C.POINTER(_i64), C.POINTER(_i64), _i64, _vp, _vp, _vp, _vp,
                                      _vp, _vp]),
    "ptgs_debug_sweep_stats": (C.c_int, [C.POINTER(C.c_uint64), _i32]),
    "ptgs_debug_fma_probe": (C.c_int, [_vp, _i32, _vp]),
    "ptgs_spatial_order_workspace_bytes": (_sz, [_i64]),
    "ptgs_spatial_order": (C.c_int, [_vp, _vp, _vp, _vp, _i64, _vp, _vp, _vp, _vp, _vp, _sz, _vp]),
    "ptgs_unpermute_u8": (C.c_int, [_vp, _vp, _i64, _vp, _vp]),
    "ptgs_vast_point_range": (C.c_int, [C.c_int, _vp, _vp, _vp, _vp, _vp, _i32, _vp, _vp, _sz, _vp]),
    "ptgs_vast_airspace": (C.c_int, [_vp, _vp, _vp, _vp, _i64, _vp, _vp]),
    "ptgs_vast_begin_block": (C.c_int, [_vp, _i64, _vp]),
    "ptgs_vast_coverage_workspace_bytes": (_sz, [_i64, _i32]),
    "ptgs_vast_coverage": (C.c_int, [C.c_int, _vp, _vp, _vp, _vp, _i64, _vp, _vp, _i32, _i32, _vp, _vp, _sz, _vp]),
    "ptgs_vast_mark_own": (C.c_int, [_vp, _i64, _i32, _vp, _vp]),
    "ptgs_vast_unique_workspace_bytes": (_sz, [_i64]),
    "ptgs_vast_unique": (C.c_int, [C.c_int, _vp, _vp, _vp, _i64, _vp, _vp, _i64, _vp, _vp, _sz, _vp]),
    "ptgs_coverage_workspace_bytes": (_sz, [_i64, _i32]),
    "ptgs_coverage_sweep": (C.c_int, [_vp, _vp, _vp, _i64, _vp, _i32, _i32, _vp, _vp, _vp, _sz, _vp]),
}


class PtgsError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("libptgs error %d: %s" % (code, msg))
        self.code = code


_lib = None


def load(build_if_missing: bool = True):
    """Loads libptgs.so (building it with nvcc when absent).  Raises when the
    CUDA library cannot be had -- there is no fallback path."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        if not build_if_missing:
            raise ImportError("libptgs.so not built (run __graft_entry__.build())")
        from . import build as _build
        _build.build()
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the symbol is missing
        fn.restype = res
        fn.argtypes = args
    global EXTREME_DIRS, POLY_FLOATS
    EXTREME_DIRS = int(lib.ptgs_extreme_dirs())
    POLY_FLOATS = 4 + 3 * EXTREME_DIRS
    _lib = lib
    return lib


def check(code: int):
    if code != PTGS_OK:
        raise PtgsError(code, load().ptgs_last_error().decode("utf-8", "replace"))
