"""Rest of Frame construction on the GPU (orbx_undistort_points / orbx_frame_finish through the C ABI)
against the committed outputs of cv2 4.13 and of the reference's own src/Frame.cc
(tests/golden/frame_finish.npz), and against the restatement on seeded random scenes.  Bit-exact."""
import os
import sys

import numpy as np
import pytest

import oracle_lib
from oracle_lib import oracle

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
from make_golden_finish_cases import CAMERAS, FINISH_CASES, depth_image, keypoints, random_points  # noqa: E402

pytestmark = pytest.mark.gpu
REF = np.load(os.path.join(HERE, "golden", "frame_finish.npz"))
DISTORTED = sorted(k for k, v in CAMERAS.items() if v[3][0] != 0.0)


def _cam(name):
    from orb_slam2_e_b200 import frame
    w, h, cam, dist, bf = CAMERAS[name]
    d = list(dist) + [0.0] * (5 - len(dist))
    return frame.camera(*cam, *d, bf=bf)


def _keys(x, y):
    from orb_slam2_e_b200 import _lib
    k = np.zeros(len(x), _lib.KEYPOINT_DTYPE)
    k["x"], k["y"], k["size"] = x, y, 31.0
    return k


def _finish(camname, x, y, depth, **kw):
    from orb_slam2_e_b200 import frame
    w, h = CAMERAS[camname][:2]
    cam = _cam(camname)
    b = frame.image_bounds(w, h, cam)
    fb = frame.frame_bounds(*b)
    out = frame.frame_finish(_keys(x, y), cam, fb, depth, **kw)
    out["bounds"] = np.array([b[0], b[1], b[2], b[3], fb.width_inv, fb.height_inv], np.float32)
    out["cell_idx"] = out["cell_idx"].astype(np.int32)
    return out


@pytest.mark.parametrize("name", DISTORTED)
def test_undistort_points_equals_cv2_fixture(name):
    from orb_slam2_e_b200 import frame
    got = frame.undistort_points(random_points(name, 4000, 5), _cam(name))
    assert got.tobytes() == REF["cv2/" + name].tobytes()


@pytest.mark.parametrize("case", sorted(FINISH_CASES))
def test_frame_finish_reproduces_reference_fixture(case):
    camname, n, seed, with_depth = FINISH_CASES[case]
    x, y = keypoints(camname, n, seed)
    got = _finish(camname, x, y, depth_image(camname, seed) if with_depth else None)
    for k in ("bounds", "xy_un", "cell_off", "cell_idx", "u_right", "depth"):
        key = "ref/%s/%s" % (case, k)
        if key in REF.files:
            assert got[k].tobytes() == REF[key].tobytes(), (case, k)
    assert got["n_in_grid"] == int(REF["ref/%s/cell_off" % case][-1])


@pytest.mark.parametrize("seed", range(6))
def test_frame_finish_equals_restatement_on_random_scenes(seed):
    rng = np.random.default_rng(500 + seed)
    camname = sorted(CAMERAS)[seed % len(CAMERAS)]
    n = int(rng.integers(1, 6000))
    x, y = keypoints(camname, n, 900 + seed)
    if seed % 2:  # crowd half of the keypoints into a handful of cells (long cell lists)
        x[n // 2:] = x[n // 2] + rng.integers(0, 12, n - n // 2).astype(np.float32)
        y[n // 2:] = y[n // 2] + rng.integers(0, 12, n - n // 2).astype(np.float32)
        w, h = CAMERAS[camname][:2]
        x, y = np.minimum(x, np.float32(w - 1)), np.minimum(y, np.float32(h - 1))
    depth = depth_image(camname, seed) if seed % 3 else None
    want = oracle_lib.port_frame_finish(oracle(), x, y, CAMERAS[camname], depth)
    got = _finish(camname, x, y, depth)
    for k, v in want.items():
        assert got[k].tobytes() == v.tobytes(), (seed, k)


def test_device_resident_depth_and_pitched_rows():
    import torch
    camname = "tum1"
    x, y = keypoints(camname, 1500, 3)
    depth = depth_image(camname, 3)
    want = _finish(camname, x, y, depth)
    h, w = depth.shape
    pitched = torch.zeros((h, w + 24), dtype=torch.float32, device="cuda")
    pitched[:, :w] = torch.from_numpy(depth).cuda()
    torch.cuda.synchronize()
    got = _finish(camname, x, y, None, depth_device_ptr=pitched.data_ptr(), depth_shape=(h, w), depth_stride=4 * (w + 24))
    for k in ("xy_un", "u_right", "depth", "cell_off", "cell_idx"):
        assert got[k].tobytes() == want[k].tobytes(), k


def test_empty_and_bad_arguments():
    import ctypes as C
    from orb_slam2_e_b200 import _lib, frame
    cam = _cam("tum1")
    fb = frame.frame_bounds(0, 640, 0, 480)
    out = frame.frame_finish(np.zeros(0, _lib.KEYPOINT_DTYPE), cam, fb)
    assert out["n_in_grid"] == 0 and not out["cell_off"].any() and out["xy_un"].shape == (0, 2)
    assert frame.undistort_points(np.zeros((0, 2), np.float32), cam).shape == (0, 2)
    with pytest.raises(_lib.OrbFrontError) as e:
        frame.frame_finish(np.zeros(4, _lib.KEYPOINT_DTYPE), cam, _lib.FrameBounds(0, 0, 0.1, 0.1, 0, 48))
    assert e.value.code == _lib.ORB_E_INVALID
    # a non-default grid (KeyFrame-style cols x rows) keeps the (ix, iy) -> ix*rows+iy order
    x, y = keypoints("tum3", 500, 8)
    fb = frame.frame_bounds(0, 640, 0, 480, cols=10, rows=7)
    got = frame.frame_finish(_keys(x, y), _cam("tum3"), fb)
    off = np.zeros(71, np.int32); idx = np.zeros(500, np.int32)
    xy = np.zeros((500, 2), np.float32)
    tot = oracle().orc_frame_finish(np.ascontiguousarray(np.stack([x, y], 1)), 500, np.array(CAMERAS["tum3"][2], np.float32),
                                    np.zeros(4, np.float32), 4, 0.0, fb.min_x, fb.min_y, fb.width_inv, fb.height_inv, 10, 7, None, 640,
                                    xy, np.zeros(500, np.float32), np.zeros(500, np.float32), off, idx)
    assert np.array_equal(got["cell_off"], off) and np.array_equal(got["cell_idx"].astype(np.int32), idx[:tot])


@pytest.mark.parametrize("case", ["tum1_rgbd", "euroc_mono", "tum3_rgbd_nodist"])
def test_cpp_finish_frame_reproduces_reference_fixture(case):
    """include/ORBextractor.h FinishFrame / ComputeImageBounds (the calls a patched Frame.cc makes) over cv:: types."""
    import dropin_lib
    L = dropin_lib.load()
    camname, n, seed, with_depth = FINISH_CASES[case]
    w, h, cam, dist, bf = CAMERAS[camname]
    x, y = keypoints(camname, n, seed)
    k = _keys(x, y)
    k["octave"] = np.arange(n) % 8
    k["angle"] = np.linspace(0, 359, n, dtype=np.float32)
    depth = depth_image(camname, seed) if with_depth else None
    bounds = np.zeros(6, np.float32); xy = np.zeros((n, 2), np.float32)
    ur, dz = np.zeros(n, np.float32), np.zeros(n, np.float32)
    off = np.zeros(64 * 48 + 1, np.int32); idx = np.zeros(n, np.int32)
    tot = L.dropin_finish_frame(k.ctypes.data, n, np.array(cam, np.float32), np.array(dist, np.float32), len(dist), bf, w, h,
                                None if depth is None else depth.ctypes.data, bounds, xy, ur, dz, off, idx)
    assert tot >= 0, L.dropin_error()
    got = dict(bounds=bounds, xy_un=xy, cell_off=off, cell_idx=idx[:tot], u_right=ur, depth=dz)
    for name in ("bounds", "xy_un", "cell_off", "cell_idx", "u_right", "depth"):
        key = "ref/%s/%s" % (case, name)
        if key in REF.files:
            assert got[name].tobytes() == REF[key].tobytes(), (case, name)
