"""ctypes bindings to the CPU oracle (oracle/_build/liborboracle.so) and, when present, to the
reference build (oracle/_ref/liborbref.so).  TEST INFRASTRUCTURE: imported only from tests/,
__graft_entry__.smoke() and bench.py's CPU-baseline legs."""
import ctypes as C
import os
import subprocess
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
u8p = np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")
i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")
i64p = np.ctypeslib.ndpointer(np.int64, flags="C_CONTIGUOUS")
u32p = np.ctypeslib.ndpointer(np.uint32, flags="C_CONTIGUOUS")
f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")

KP_DTYPE = np.dtype([("x", "f4"), ("y", "f4"), ("size", "f4"), ("angle", "f4"), ("response", "f4"),
                     ("octave", "i4"), ("class_id", "i4")])


def build_oracle(fast=False):
    name = "liborboracle_fast.so" if fast else "liborboracle.so"
    path = os.path.join(ORACLE_DIR, "_build", name)
    srcs = [os.path.join(ORACLE_DIR, f) for f in ("oracle_capi.cc", "cvprims.hpp", "orb_port.hpp",
                                                  "orb_port_impl.hpp", "match_port.hpp")]
    if not os.path.exists(path) or any(os.path.getmtime(s) > os.path.getmtime(path) for s in srcs):
        subprocess.check_call(["make", "-C", ORACLE_DIR, "_build/" + name], stdout=subprocess.DEVNULL)
    return path


_cache = {}


def oracle(fast=False):
    key = ("port", fast)
    if key in _cache:
        return _cache[key]
    L = C.CDLL(build_oracle(fast))
    L.orc_resize_linear.argtypes = [u8p, C.c_int, C.c_int, C.c_size_t, u8p, C.c_int, C.c_int, C.c_size_t]
    L.orc_border101.argtypes = [u8p, C.c_int, C.c_int, C.c_size_t, u8p, C.c_size_t, C.c_int]
    L.orc_blur7.argtypes = [u8p, C.c_int, C.c_int, C.c_size_t, u8p, C.c_size_t]
    L.orc_fast_atan2.argtypes = [C.c_float, C.c_float]
    L.orc_fast_atan2.restype = C.c_float
    L.orc_fast_atan2_v.argtypes = [f32p, f32p, f32p, C.c_int]
    L.orc_fast.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_int, i32p, C.c_int]
    L.orc_fast_strength_map.argtypes = [u8p, C.c_int, C.c_int, C.c_size_t, u8p]
    L.orc_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
    L.orc_create.restype = C.c_void_p
    L.orc_destroy.argtypes = [C.c_void_p]
    for f in (L.orc_quota, L.orc_umax):
        f.argtypes = [C.c_void_p, C.c_int]
    L.orc_scale.argtypes = [C.c_void_p, C.c_int]
    L.orc_scale.restype = C.c_float
    L.orc_extract.argtypes = [C.c_void_p, u8p, C.c_int, C.c_int, C.c_size_t, C.c_void_p, u8p, C.c_int]
    L.orc_level_size.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.orc_get_level.argtypes = [C.c_void_p, C.c_int, u8p, C.c_size_t, C.c_int]
    L.orc_stage.argtypes = [C.c_void_p, C.c_int, C.c_int, i32p, C.c_int]
    L.orc_distribute.argtypes = [C.c_void_p, i32p, C.c_int, C.c_int, C.c_int, C.c_int, i32p, C.c_int]
    L.orc_extract_batch_mt.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int, u8p, C.c_int, C.c_int, C.c_int,
                                       C.c_int, C.c_void_p]
    L.orc_extract_batch_mt.restype = C.c_longlong
    L.orc_hamming.argtypes = [u8p, u8p]
    L.orc_top2.argtypes = [u8p, C.c_int, u8p, C.c_longlong, i32p, i64p, i32p]
    L.orc_top2_mt.argtypes = [u8p, C.c_int, u8p, C.c_longlong, i32p, i64p, i32p, C.c_int]
    L.orc_accept.argtypes = [C.c_int, C.c_int, C.c_float, C.c_int, C.c_int]
    L.orc_merge_top2.argtypes = [i32p, i64p, i32p, C.c_int, C.c_int, i32p, i64p, i32p]
    L.orc_search_bow.argtypes = [u8p, C.c_int, u8p, C.c_int, u32p, i32p, u32p, C.c_int, u32p, i32p, u32p, C.c_int,
                                 u8p, C.c_void_p, f32p, f32p, C.c_float, C.c_int, C.c_int, i32p]
    L.orc_stereo_search.argtypes = [f32p, f32p, i32p, u8p, C.c_int, f32p, f32p, i32p, u8p, C.c_int, f32p, C.c_int,
                                    C.c_float, C.c_float, i32p, i32p]
    L.orc_window_search.argtypes = [f32p, f32p, i32p, C.c_void_p, u8p, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float,
                                    C.c_int, C.c_int, i32p, u32p, C.c_void_p, u8p, C.c_int, C.c_void_p, C.c_int, C.c_float,
                                    C.c_int, C.c_void_p, C.c_void_p, C.c_double, C.c_double]
    L.orc_compute_stereo_matches.argtypes = [f32p, f32p, i32p, u8p, C.c_int, f32p, f32p, i32p, u8p, C.c_int, f32p, f32p, C.c_int,
                                             C.c_void_p, C.c_void_p, i32p, i32p, i32p, C.c_float, C.c_float, f32p, f32p]
    L.orc_distinctive_descriptors.argtypes = [u8p, i32p, C.c_int, i32p]
    L.orc_undistort_points.argtypes = [f32p, C.c_int, f32p, f32p, C.c_int, f32p]
    L.orc_image_bounds.argtypes = [C.c_int, C.c_int, f32p, f32p, C.c_int, f32p]
    L.orc_frame_finish.argtypes = [f32p, C.c_int, f32p, f32p, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, C.c_float,
                                   C.c_int, C.c_int, C.c_void_p, C.c_int, f32p, f32p, f32p, i32p, i32p]
    L.orc_features_in_area.argtypes = [f32p, f32p, i32p, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, C.c_int, C.c_int,
                                       i32p, u32p, f32p, f32p, f32p, i32p, i32p, C.c_int, i32p, i32p, C.c_int]
    _cache[key] = L
    return L


def ref_lib(fast=False):
    """The reference's own ORBextractor.cc over the shim; None when not built (GPU box builds
    nothing from /root/reference, but a prebuilt oracle/_ref travels with the snapshot)."""
    key = ("ref", fast)
    if key in _cache:
        return _cache[key]
    path = os.path.join(ORACLE_DIR, "_ref", "liborbref_fast.so" if fast else "liborbref.so")
    if not os.path.exists(path):
        if os.path.isdir("/root/reference/src"):
            subprocess.check_call(["make", "-C", ORACLE_DIR, "ref"], stdout=subprocess.DEVNULL)
        else:
            _cache[key] = None
            return None
    L = C.CDLL(path)
    L.orbref_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
    L.orbref_create.restype = C.c_void_p
    L.orbref_destroy.argtypes = [C.c_void_p]
    L.orbref_extract.argtypes = [C.c_void_p, u8p, C.c_int, C.c_int, C.c_size_t, C.c_void_p, u8p, C.c_int]
    L.orbref_stage_keypoints.argtypes = [C.c_void_p, u8p, C.c_int, C.c_int, C.c_size_t, C.c_void_p, i32p, C.c_int]
    L.orbref_level_size.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.orbref_get_level.argtypes = [C.c_void_p, C.c_int, u8p, C.c_size_t, C.c_int]
    L.orbref_quota.argtypes = [C.c_void_p, C.c_int]
    L.orbref_umax.argtypes = [C.c_void_p, C.c_int]
    L.orbref_scale.argtypes = [C.c_void_p, C.c_int]
    L.orbref_scale.restype = C.c_float
    _cache[key] = L
    return L


def frame_ref_lib():
    """The reference's own Frame.cc + MapPoint.cc (+ ORBextractor.cc, DBoW2 vectors) over the shim
    (oracle/_ref/libframeref.so, which imports ORBmatcher's constants and DescriptorDistance from
    libmatchref.so); None when it is not built and cannot be (no /root/reference)."""
    key = ("frameref",)
    if key in _cache:
        return _cache[key]
    path = os.path.join(ORACLE_DIR, "_ref", "libframeref.so")
    if not os.path.exists(path):
        if os.path.isdir("/root/reference/src"):
            subprocess.check_call(["make", "-C", ORACLE_DIR, "_ref/libframeref.so"], stdout=subprocess.DEVNULL)
        else:
            _cache[key] = None
            return None
    L = C.CDLL(path)
    L.frameref_constants.argtypes = [i32p]
    L.frameref_stereo.argtypes = [C.c_int, C.c_float, C.c_int, f32p, f32p, i32p, u8p, C.c_int, f32p, f32p, i32p, u8p, C.c_int,
                                  C.c_void_p, C.c_void_p, i32p, i32p, i32p, C.c_float, C.c_float, f32p, f32p]
    L.frameref_features_in_area.argtypes = [f32p, f32p, i32p, C.c_int, C.c_float, C.c_float, C.c_float, C.c_float, f32p, f32p,
                                            f32p, i32p, i32p, C.c_int, i32p, i32p, C.c_int]
    L.frameref_distinctive.argtypes = [u8p, i32p, C.c_void_p, C.c_int, u8p]
    L.frameref_finish.argtypes = [f32p, f32p, C.c_int, f32p, f32p, C.c_int, C.c_float, C.c_int, C.c_int, C.c_void_p, f32p, f32p,
                                  f32p, f32p, i32p, i32p]
    _cache[key] = L
    return L


def ref_frame_finish(L, x, y, camera, depth=None):
    """The reference's own Frame code (libframeref.so) after extraction: camera = (w, h, (fx, fy, cx, cy), dist, bf)."""
    w, h, cam, dist, bf = camera
    n = len(x)
    cam = np.array(cam, np.float32)
    dist = np.array(dist, np.float32)
    bounds = np.zeros(6, np.float32)
    xy = np.zeros((n, 2), np.float32)
    ur, dz = np.full(n, -2, np.float32), np.full(n, -2, np.float32)
    off = np.zeros(64 * 48 + 1, np.int32)
    idx = np.zeros(max(n, 1), np.int32)
    if depth is not None:
        assert depth.shape == (h, w) and depth.dtype == np.float32 and depth.flags.c_contiguous
    total = L.frameref_finish(np.ascontiguousarray(x, np.float32), np.ascontiguousarray(y, np.float32), n, cam, dist, len(dist), bf, w,
                              h, None if depth is None else depth.ctypes.data, bounds, xy, ur, dz, off, idx)
    out = dict(bounds=bounds, xy_un=xy, cell_off=off, cell_idx=idx[:total].copy())
    if depth is not None:
        out.update(u_right=ur, depth=dz)
    return out


def port_frame_finish(L, x, y, camera, depth=None):
    """The restatement (oracle/match_port.hpp) of the same: bounds, undistortion, RGB-D stereo, grid."""
    w, h, cam, dist, bf = camera
    n = len(x)
    cam = np.array(cam, np.float32)
    dist = np.array(dist, np.float32)
    b = np.zeros(4, np.float32)
    L.orc_image_bounds(w, h, cam, dist, len(dist), b)
    f32 = np.float32
    winv, hinv = f32(64) / (b[1] - b[0]), f32(48) / (b[3] - b[2])
    xy_in = np.ascontiguousarray(np.stack([x, y], 1), np.float32)
    xy = np.zeros((n, 2), np.float32)
    ur, dz = np.full(n, -2, np.float32), np.full(n, -2, np.float32)
    off = np.zeros(64 * 48 + 1, np.int32)
    idx = np.zeros(max(n, 1), np.int32)
    total = L.orc_frame_finish(xy_in, n, cam, dist, len(dist), bf, b[0], b[2], winv, hinv, 64, 48,
                               None if depth is None else depth.ctypes.data, w, xy, ur, dz, off, idx)
    out = dict(bounds=np.array([b[0], b[1], b[2], b[3], winv, hinv], np.float32), xy_un=xy, cell_off=off, cell_idx=idx[:total].copy())
    if depth is not None:
        out.update(u_right=ur, depth=dz)
    return out


def ref_stereo_matches(L, cfg, kl, dl, kr, dr, levels_left, levels_right, mb, mbf):
    """Frame::ComputeStereoMatches of the reference on given keypoints and pyramid levels."""
    n_levels = len(levels_left)
    PL = (C.c_void_p * n_levels)(*[a.ctypes.data for a in levels_left])
    PR = (C.c_void_p * n_levels)(*[a.ctypes.data for a in levels_right])
    for a, b in zip(levels_left, levels_right):
        assert a.shape == b.shape and a.strides == b.strides and a.strides[1] == 1
    step = np.array([a.strides[0] for a in levels_left], np.int32)
    w = np.array([a.shape[1] for a in levels_left], np.int32)
    h = np.array([a.shape[0] for a in levels_left], np.int32)
    f = lambda a: np.ascontiguousarray(a, np.float32)
    ur, z = np.zeros(len(kl), np.float32), np.zeros(len(kl), np.float32)
    L.frameref_stereo(cfg[0], cfg[1], cfg[2], f(kl["x"]), f(kl["y"]), np.ascontiguousarray(kl["octave"], np.int32),
                      np.ascontiguousarray(dl), len(kl), f(kr["x"]), f(kr["y"]), np.ascontiguousarray(kr["octave"], np.int32),
                      np.ascontiguousarray(dr), len(kr), C.cast(PL, C.c_void_p), C.cast(PR, C.c_void_p), step, w, h, mb, mbf, ur, z)
    return ur, z


class _Ext:
    """Common wrapper: extract(img) -> (kps structured array, desc (n,32) u8)."""

    def __init__(self, lib, prefix, nf, sf, nl, ini, mn):
        self.L, self.p = lib, prefix
        self.nl = nl
        self.h = getattr(lib, prefix + "_create")(nf, sf, nl, ini, mn)
        self.cap = nf + 4 * nl + 64

    def __del__(self):
        try:
            getattr(self.L, self.p + "_destroy")(self.h)
        except Exception:
            pass

    def extract(self, img):
        img = np.ascontiguousarray(img, np.uint8)
        kps = np.zeros(self.cap, KP_DTYPE)
        desc = np.zeros((self.cap, 32), np.uint8)
        n = getattr(self.L, self.p + "_extract")(self.h, img, img.shape[1], img.shape[0], img.strides[0],
                                                  kps.ctypes.data, desc, self.cap)
        assert n >= 0, n
        return kps[:n].copy(), desc[:n].copy()

    def level(self, l, with_border=False):
        w, h = C.c_int(), C.c_int()
        assert getattr(self.L, self.p + "_level_size")(self.h, l, C.byref(w), C.byref(h)) == 0
        b = 19 if with_border else 0
        out = np.zeros((h.value + 2 * b, w.value + 2 * b), np.uint8)
        getattr(self.L, self.p + "_get_level")(self.h, l, out, out.strides[0], int(with_border))
        return out

    def quota(self, l):
        return getattr(self.L, self.p + "_quota")(self.h, l)

    def umax(self, v):
        return getattr(self.L, self.p + "_umax")(self.h, v)

    def scale(self, l):
        return getattr(self.L, self.p + "_scale")(self.h, l)


class PortExtractor(_Ext):
    def __init__(self, nf, sf, nl, ini, mn, fast=False):
        super().__init__(oracle(fast), "orc", nf, sf, nl, ini, mn)

    def stage(self, which, l, cap=1 << 20):
        buf = np.zeros((cap, 3), np.int32)
        n = self.L.orc_stage(self.h, which, l, buf.reshape(-1), cap)
        return buf[:n].copy()

    def distribute(self, cand, box_w, box_h, N):
        cand = np.ascontiguousarray(cand, np.int32)
        out = np.zeros((N + 8, 3), np.int32)
        n = self.L.orc_distribute(self.h, cand.reshape(-1), len(cand), box_w, box_h, N, out.reshape(-1), N + 8)
        return out[:n].copy()


class RefExtractor(_Ext):
    def __init__(self, nf, sf, nl, ini, mn, fast=False):
        lib = ref_lib(fast)
        assert lib is not None
        super().__init__(lib, "orbref", nf, sf, nl, ini, mn)
