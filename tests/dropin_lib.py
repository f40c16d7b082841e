"""ctypes access to tests/cpp/_build/libdropin.so: the product's C++ host classes
(include/ORBextractor.h, include/ORBmatcher.h) driven the way the reference's callers drive them."""
import ctypes as C
import os
import subprocess

import numpy as np
from numpy.ctypeslib import ndpointer

HERE = os.path.dirname(os.path.abspath(__file__))
CPP = os.path.join(HERE, "cpp")
PATH = os.path.join(CPP, "_build", "libdropin.so")
u8p = ndpointer(np.uint8, flags="C_CONTIGUOUS")
i32p = ndpointer(np.int32, flags="C_CONTIGUOUS")
u32p = ndpointer(np.uint32, flags="C_CONTIGUOUS")
f32p = ndpointer(np.float32, flags="C_CONTIGUOUS")

_lib = None


def build():
    from orb_slam2_e_b200 import _lib as core
    core.load()  # liborbfront.so must exist before linking against it
    subprocess.check_call(["make", "-s", "-C", CPP], stdout=subprocess.DEVNULL)
    return PATH


def load():
    global _lib
    if _lib is not None:
        return _lib
    build()
    L = C.CDLL(PATH)
    L.dropin_error.restype = C.c_char_p
    L.dropin_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
    L.dropin_create.restype = C.c_void_p
    L.dropin_destroy.argtypes = [C.c_void_p]
    L.dropin_extract.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_void_p, u8p, C.c_int]
    L.dropin_time_extract.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_int, C.c_int]
    L.dropin_time_extract.restype = C.c_double
    L.dropin_level_size.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.dropin_get_level.argtypes = [C.c_void_p, C.c_int, u8p, C.c_size_t, C.c_int]
    L.dropin_scale.argtypes = [C.c_void_p, C.c_int]
    L.dropin_scale.restype = C.c_float
    L.dropin_table.argtypes = [C.c_void_p, C.c_int, C.c_int]
    L.dropin_table.restype = C.c_float
    L.dropin_levels.argtypes = [C.c_void_p]
    L.dropin_scale_factor.argtypes = [C.c_void_p]
    L.dropin_scale_factor.restype = C.c_float
    L.dropin_distance.argtypes = [u8p, u8p]
    L.dropin_search_bow.argtypes = [u8p, C.c_int, u8p, C.c_int, u32p, i32p, u32p, C.c_int, u32p, i32p, u32p, C.c_int,
                                    u8p, u8p, f32p, f32p, C.c_float, C.c_int, C.c_int, i32p]
    pp = C.POINTER(C.c_void_p)
    L.dropin_search_bow_batch.argtypes = [C.c_int, pp, i32p, pp, pp, pp, i32p, pp, pp, u8p, C.c_int, u32p, i32p, u32p, C.c_int, f32p,
                                          C.c_float, C.c_int, pp, i32p]
    L.dropin_stereo.argtypes = [C.c_void_p, u8p, C.c_int, C.c_void_p, u8p, C.c_int, f32p, C.c_int, C.c_int, C.c_float,
                                C.c_float, i32p, i32p]
    L.dropin_stereo_frame.argtypes = [C.c_void_p, C.c_void_p, u8p, u8p, C.c_int, C.c_int, C.c_size_t, C.c_float, C.c_float,
                                      C.c_void_p, u8p, f32p, f32p, C.c_int]
    L.dropin_finish_frame.argtypes = [C.c_void_p, C.c_int, f32p, f32p, C.c_int, C.c_float, C.c_int, C.c_int, C.c_void_p, f32p, f32p,
                                      f32p, f32p, i32p, i32p]
    _lib = L
    return L


SYMBOLS = ["dropin_error", "dropin_create", "dropin_destroy", "dropin_extract", "dropin_time_extract", "dropin_level_size", "dropin_get_level",
           "dropin_scale", "dropin_table", "dropin_levels", "dropin_scale_factor", "dropin_distance", "dropin_search_bow", "dropin_search_bow_batch",
           "dropin_stereo", "dropin_stereo_frame", "dropin_finish_frame"]
