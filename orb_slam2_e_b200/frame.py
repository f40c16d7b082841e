"""Host-side mirror of the rest of Frame construction (include/orbfront.h):
Frame::UndistortKeyPoints / ComputeImageBounds / AssignFeaturesToGrid / ComputeStereoFromRGBD
(src/Frame.cc:419-485, :245-260, :703-727) over orbx_undistort_points and orbx_frame_finish.
No CPU fallback: without a CUDA device every call raises OrbFrontError(ORB_E_NO_DEVICE)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import Camera, FrameBounds, KEYPOINT_DTYPE, check, ptr

FRAME_GRID_COLS, FRAME_GRID_ROWS = 64, 48  # include/Frame.h:39-40


def camera(fx, fy, cx, cy, k1=0.0, k2=0.0, p1=0.0, p2=0.0, k3=0.0, bf=0.0) -> Camera:
    return Camera(fx, fy, cx, cy, k1, k2, p1, p2, k3, bf)


def undistort_points(xy, cam: Camera, device: int = 0) -> np.ndarray:
    """cv::undistortPoints(xy, xy, K, distCoef, Mat(), K) as Frame.cc:438 / :466 call it."""
    p = np.ascontiguousarray(xy, np.float32).reshape(-1, 2)
    out = np.empty_like(p)
    check(_lib.load().orbx_undistort_points(device, ptr(p) if len(p) else None, len(p), C.byref(cam),
                                            ptr(out) if len(p) else None))
    return out


def image_bounds(width: int, height: int, cam: Camera, device: int = 0):
    """Frame::ComputeImageBounds (Frame.cc:451-485): (mnMinX, mnMaxX, mnMinY, mnMaxY) as float32."""
    f32 = np.float32
    if cam.k1 == 0.0:
        return f32(0.0), f32(width), f32(0.0), f32(height)
    m = undistort_points([[0.0, 0.0], [width, 0.0], [0.0, height], [width, height]], cam, device)
    return min(m[0, 0], m[2, 0]), max(m[1, 0], m[3, 0]), min(m[0, 1], m[1, 1]), max(m[2, 1], m[3, 1])


def frame_bounds(min_x, max_x, min_y, max_y, cols=FRAME_GRID_COLS, rows=FRAME_GRID_ROWS) -> FrameBounds:
    """mfGridElementWidthInv / HeightInv as Frame.cc:106-107 computes them (float arithmetic)."""
    f32 = np.float32
    return FrameBounds(f32(min_x), f32(min_y), f32(cols) / (f32(max_x) - f32(min_x)), f32(rows) / (f32(max_y) - f32(min_y)),
                       cols, rows)


def frame_finish(keys, cam: Camera, bounds: FrameBounds, depth=None, device: int = 0, depth_device_ptr=None,
                 depth_shape=None, depth_stride=None):
    """orbx_frame_finish: returns dict(xy_un (n,2) f32, cell_off (cols*rows+1) i32, cell_idx u32 [n_in_grid],
    and with a depth image u_right / depth (n) f32).  depth: (h, w) float32 array in host memory, or
    depth_device_ptr + depth_shape (+ depth_stride in bytes) for an image already on the device."""
    k = np.ascontiguousarray(keys, KEYPOINT_DTYPE)
    n = len(k)
    ncells = bounds.cols * bounds.rows
    xy = np.empty((n, 2), np.float32)
    off = np.zeros(ncells + 1, np.int32)
    idx = np.zeros(max(n, 1), np.uint32)
    ur = dz = None
    dptr, dw, dh, dstride, dmem = None, 0, 0, 0, 0
    if depth is not None:
        d = np.ascontiguousarray(depth, np.float32)
        dptr, (dh, dw), dstride = ptr(d), d.shape, d.strides[0]
    elif depth_device_ptr is not None:
        dptr, (dh, dw), dmem = C.c_void_p(depth_device_ptr), depth_shape, 1
        dstride = depth_stride if depth_stride else 4 * dw
    if dptr is not None:
        ur = np.empty(n, np.float32)
        dz = np.empty(n, np.float32)
    total = C.c_int(0)
    check(_lib.load().orbx_frame_finish(device, ptr(k) if n else None, n, C.byref(cam), C.byref(bounds), dptr, dw, dh, dstride, dmem,
                                        ptr(xy) if n else None, None if ur is None else ptr(ur), None if dz is None else ptr(dz),
                                        ptr(off), ptr(idx), C.byref(total)))
    out = dict(xy_un=xy, cell_off=off, cell_idx=idx[: total.value], n_in_grid=total.value)
    if ur is not None:
        out.update(u_right=ur, depth=dz)
    return out
