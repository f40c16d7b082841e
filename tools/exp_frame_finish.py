"""Latency of the rest-of-Frame call (orbx_frame_finish, host buffers in and out) against the
restatement of the reference's Frame code on one host core."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests")); sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
import numpy as np, torch
import oracle_lib
from make_golden_finish_cases import CAMERAS, depth_image, keypoints
from orb_slam2_e_b200 import _lib, frame

def t(fn, reps=200):
    for _ in range(10): fn()
    t0 = time.perf_counter()
    for _ in range(reps): fn()
    return (time.perf_counter() - t0) / reps * 1e6

O = oracle_lib.oracle(fast=True)
for camname, n in (("tum1", 1000), ("tum1", 2000), ("euroc", 8000)):
    w, h, cam, dist, bf = CAMERAS[camname]
    x, y = keypoints(camname, n, 1)
    k = np.zeros(n, _lib.KEYPOINT_DTYPE); k["x"], k["y"] = x, y
    d = list(dist) + [0.0] * (5 - len(dist))
    c = frame.camera(*cam, *d, bf=bf)
    fb = frame.frame_bounds(*frame.image_bounds(w, h, c))
    depth = depth_image(camname, 1)
    dd = torch.from_numpy(depth).cuda(); torch.cuda.synchronize()
    a = t(lambda: frame.frame_finish(k, c, fb))
    b = t(lambda: frame.frame_finish(k, c, fb, depth))
    e = t(lambda: frame.frame_finish(k, c, fb, None, depth_device_ptr=dd.data_ptr(), depth_shape=depth.shape))
    p = t(lambda: oracle_lib.port_frame_finish(O, x, y, CAMERAS[camname], depth))
    print("%s n=%d: GPU call %.0f us (mono), %.0f us (host depth %dx%d uploaded), %.0f us (device depth); restatement, one core %.0f us"
          % (camname, n, a, b, w, h, e, p))
