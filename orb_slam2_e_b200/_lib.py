"""ctypes binding of liborbfront.so (include/orbfront.h).  Fails loudly when the library is
missing or a call reports an error; there is no CPU fallback anywhere in this package."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import build as _build

ORB_OK, ORB_E_INVALID, ORB_E_NO_DEVICE, ORB_E_CUDA, ORB_E_CAPACITY, ORB_E_STATE = 0, -1, -2, -3, -4, -5

KEYPOINT_DTYPE = np.dtype([("x", "f4"), ("y", "f4"), ("size", "f4"), ("angle", "f4"), ("response", "f4"),
                           ("octave", "i4"), ("class_id", "i4")])
TOP2_DTYPE = np.dtype([("best1", "i4"), ("best2", "i4"), ("idx", "i8")])
WINDOW_QUERY_DTYPE = np.dtype([("x", "f4"), ("y", "f4"), ("r", "f4"), ("u_right", "f4"), ("er_max", "f4"),
                               ("min_level", "i4"), ("max_level", "i4"), ("blocks", "i4")])
WINDOW_RESULT_DTYPE = np.dtype([("idx", "i4"), ("best_idx", "i4"), ("best_dist", "i4"), ("best_dist2", "i4"),
                                ("best_level", "i4"), ("best_level2", "i4")])


class OrbFrontError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"liborbfront error {code}: {msg}")
        self.code = code


class OrbxConfig(C.Structure):
    _fields_ = [("nfeatures", C.c_int), ("scale_factor", C.c_float), ("nlevels", C.c_int),
                ("ini_th_fast", C.c_int), ("min_th_fast", C.c_int), ("device", C.c_int),
                ("max_width", C.c_int), ("max_height", C.c_int), ("max_batch", C.c_int)]


class FeatVec(C.Structure):
    _fields_ = [("node", C.c_void_p), ("off", C.c_void_p), ("idx", C.c_void_p), ("n_nodes", C.c_int32)]


class BowSide(C.Structure):  # orbm_bow_side
    _fields_ = [("desc", C.c_void_p), ("n", C.c_int32), ("fv", C.POINTER(FeatVec)), ("valid", C.c_void_p), ("angle", C.c_void_p)]


class GridFrame(C.Structure):  # orbm_grid_frame
    _fields_ = [("x", C.c_void_p), ("y", C.c_void_p), ("octave", C.c_void_p), ("u_right", C.c_void_p),
                ("desc", C.c_void_p), ("n", C.c_int32), ("min_x", C.c_float), ("min_y", C.c_float),
                ("width_inv", C.c_float), ("height_inv", C.c_float), ("cols", C.c_int32), ("rows", C.c_int32),
                ("cell_off", C.c_void_p), ("cell_idx", C.c_void_p)]


class TriKeyFrame(C.Structure):  # orbm_tri_keyframe
    _fields_ = [("desc", C.c_void_p), ("x", C.c_void_p), ("y", C.c_void_p), ("octave", C.c_void_p), ("angle", C.c_void_p),
                ("usable", C.c_void_p), ("stereo", C.c_void_p), ("n", C.c_int32), ("fv", C.POINTER(FeatVec))]


class WindowParams(C.Structure):  # orbm_window_params
    _fields_ = [("th", C.c_int32), ("ratio", C.c_float), ("claim", C.c_int32), ("inv_level_sigma2", C.c_void_p),
                ("n_levels", C.c_int32), ("chi2_mono", C.c_double), ("chi2_stereo", C.c_double)]


_lib = None

_vp, _i, _f, _sz, _i64 = C.c_void_p, C.c_int, C.c_float, C.c_size_t, C.c_int64
_ip = C.POINTER(C.c_int)

class Camera(C.Structure):
    _fields_ = [("fx", _f), ("fy", _f), ("cx", _f), ("cy", _f), ("k1", _f), ("k2", _f), ("p1", _f), ("p2", _f), ("k3", _f),
                ("bf", _f)]


class FrameBounds(C.Structure):
    _fields_ = [("min_x", _f), ("min_y", _f), ("width_inv", _f), ("height_inv", _f), ("cols", C.c_int32), ("rows", C.c_int32)]


# name -> (restype, argtypes); every symbol include/orbfront.h declares
SIGNATURES = {
    "orb_last_error": (C.c_char_p, []),
    "orb_device_count": (_i, []),
    "orb_launch_count": (C.c_longlong, []),
    "orbx_create": (_i, [C.POINTER(OrbxConfig), C.POINTER(_vp)]),
    "orbx_destroy": (_i, [_vp]),
    "orbx_get_scale_table": (_i, [_vp, _i, _vp]),
    "orbx_get_level_quota": (_i, [_vp, _vp]),
    "orbx_extract": (_i, [_vp, _vp, _i, _i, _sz, _vp, _vp, _i, _ip]),
    "orbx_extract_batch": (_i, [_vp, _vp, _i, _i, _i, _sz, _sz, _i]),
    "orbx_extract_batch_host": (_i, [_vp, _vp, _i, _i, _i, _sz, _sz, _vp, _vp, _i, _vp, _i]),
    "orbx_frame_capacity": (_i, [_vp, _i, _i, _ip]),
    "orbx_batch_counts": (_i, [_vp, _vp]),
    "orbx_batch_read": (_i, [_vp, _i, _vp, _vp, _i, _ip]),
    "orbx_batch_read_all": (_i, [_vp, _vp, _vp, _i, _vp]),
    "orbx_batch_read_begin": (_i, [_vp, _vp, _vp, _i, _vp]),
    "orbx_batch_read_end": (_i, [_vp]),
    "orbx_sync": (_i, [_vp]),
    "orbx_level_size": (_i, [_vp, _i, _ip, _ip]),
    "orbx_get_level": (_i, [_vp, _i, _i, _vp, _sz, _i]),
    "orbx_set_pyramid_staging": (_i, [_vp, _i]),
    "orbx_get_pyramid_host": (_i, [_vp, _i, C.POINTER(_vp), _ip, _ip, C.POINTER(_sz)]),
    "orbx_read_frame0": (_i, [_vp, _vp, _vp, _i, _ip]),
    "orbx_get_blur_level": (_i, [_vp, _i, _i, _vp, _sz]),
    "orbx_get_candidates": (_i, [_vp, _i, _i, _vp, _i, _ip]),
    "orbx_last_batch_ms": (_i, [_vp, C.POINTER(_f), _ip]),
    "orbx_get_stream": (_i, [_vp, C.POINTER(_vp)]),
    "orbx_set_stage_timing": (_i, [_vp, _i]),
    "orbx_last_stage_ms": (_i, [_vp, _vp]),
    "orbm_db_get_stream": (_i, [_vp, C.POINTER(_vp)]),
    "orbm_hamming": (_i, [_vp, _vp]),
    "orbm_hamming_pairs": (_i, [_i, _vp, _vp, _i, _vp]),
    "orbm_db_create": (_i, [_i, _i64, C.POINTER(_vp)]),
    "orbm_db_append": (_i, [_vp, _vp, _i64, _i]),
    "orbm_db_size": (_i, [_vp, C.POINTER(_i64)]),
    "orbm_db_destroy": (_i, [_vp]),
    "orbm_top2_search": (_i, [_vp, _vp, _i, _vp, _i]),
    "orbm_top2_search_noalloc": (_i, [_i, _vp, _i64, _i64, _vp, _i, _vp, _vp]),
    "orbm_db_sync": (_i, [_vp]),
    "orbm_db_last_ms": (_i, [_vp, C.POINTER(_f), _ip]),
    "orbm_merge_top2": (_i, [_vp, _i, _i, _f, _i, _vp, _vp]),
    "orbm_group_unique_id": (_i, [_vp]),
    "orbm_group_create": (_i, [_vp, _i, _i, _vp, C.POINTER(_vp)]),
    "orbm_group_create_local": (_i, [_vp, _i, C.POINTER(_vp)]),
    "orbm_group_search": (_i, [_vp, _vp, _i, _f, _i, _vp, _vp]),
    "orbm_group_last_ms": (_i, [_vp, C.POINTER(_f), C.POINTER(_f)]),
    "orbm_group_destroy": (_i, [_vp]),
    "orbm_search_bow": (_i, [_i, _vp, _i, _vp, _i, C.POINTER(FeatVec), C.POINTER(FeatVec), _vp, _vp, _vp, _vp, _f,
                             _i, _i, _vp, _ip]),
    "orbm_search_bow_batch": (_i, [_i, C.POINTER(BowSide), _i, C.POINTER(BowSide), _f, _i, _i, C.POINTER(_vp), _vp]),
    "orbm_stereo_search": (_i, [_i, _vp, _vp, _i, _vp, _vp, _i, _vp, _i, _i, _f, _f, _vp, _vp]),
    "orbm_window_search": (_i, [_i, C.POINTER(GridFrame), _vp, _vp, _i, _vp, C.POINTER(WindowParams), _vp, _ip, _ip]),
    "orbm_window_candidates": (_i, [_i, C.POINTER(GridFrame), _vp, _vp, _i, _vp, _vp, _vp, _vp, _i, _ip]),
    "orbm_search_triangulation": (_i, [_i, C.POINTER(TriKeyFrame), C.POINTER(TriKeyFrame), _vp, _f, _f, _vp, _vp, _i, _i, _vp, _ip]),
    "orbx_stereo_matches": (_i, [_vp, _i, _vp, _i, _vp, _vp, _i, _vp, _vp, _i, _f, _f, _vp, _vp]),
    "orbx_stereo_matches_batch": (_i, [_vp, _i, _f, _f, _vp, _vp, _i]),
    "orbx_undistort_points": (_i, [_i, _vp, _i, C.POINTER(Camera), _vp]),
    "orbx_frame_finish": (_i, [_i, _vp, _i, C.POINTER(Camera), C.POINTER(FrameBounds), _vp, _i, _i, _sz, _i, _vp, _vp, _vp,
                               _vp, _vp, _ip]),
    "orbm_vocab_create": (_i, [_i, _i, _i, _i, _vp, _vp, _vp, _vp, C.POINTER(C.c_void_p)]),
    "orbm_vocab_load_text": (_i, [_i, C.c_char_p, C.POINTER(C.c_void_p)]),
    "orbm_vocab_info": (_i, [_vp, _ip, _ip, _ip, _ip]),
    "orbm_vocab_destroy": (_i, [_vp]),
    "orbm_vocab_transform": (_i, [_vp, _vp, _i, _i, _vp, _vp, _vp]),
    "orbm_distinctive_descriptors": (_i, [_i, _vp, _vp, _i, _vp]),
    "orbm_db_save": (_i, [_vp, C.c_char_p]),
    "orbm_db_load": (_i, [_vp, C.c_char_p, _i64, _i64]),
    "orbm_popc_peak": (_i, [_i, C.POINTER(C.c_double)]),
}


def lib_path() -> str:
    return _build.LIB


def load():
    """Loads (building first if sources are newer) the C-ABI library."""
    global _lib
    if _lib is not None:
        return _lib
    path = _build.LIB
    if _build.needs_build():
        if os.path.exists(_build.NVCC):
            _build.build()
        elif not os.path.exists(path):
            raise OrbFrontError(ORB_E_NO_DEVICE, f"{path} is missing and nvcc is unavailable to build it")
    L = C.CDLL(path)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(L, name)  # AttributeError if the library does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


def check(rc: int):
    if rc != ORB_OK:
        raise OrbFrontError(rc, load().orb_last_error().decode(errors="replace"))


def ptr(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)
