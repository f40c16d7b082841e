"""Rest of Frame construction, CPU side: the restatement of cv::undistortPoints is
pinned on cv2 4.13 (committed outputs, and the live wheel when importable), and the restatement of
Frame::UndistortKeyPoints / ComputeImageBounds / ComputeStereoFromRGBD / AssignFeaturesToGrid on the
reference's own src/Frame.cc (tests/golden/frame_finish.npz, made by tests/golden/make_golden_finish.py;
live oracle/_ref/libframeref.so when present).  Bit-exact throughout."""
import os
import sys

import numpy as np
import pytest

import oracle_lib
from oracle_lib import oracle

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
from make_golden_finish_cases import CAMERAS, FINISH_CASES, depth_image, keypoints, random_points  # noqa: E402

REF = np.load(os.path.join(HERE, "golden", "frame_finish.npz"))
DISTORTED = sorted(k for k, v in CAMERAS.items() if v[3][0] != 0.0)


def _undistort(name, p):
    w, h, cam, dist, bf = CAMERAS[name]
    out = np.zeros_like(p)
    oracle().orc_undistort_points(p, len(p), np.array(cam, np.float32), np.array(dist, np.float32), len(dist), out)
    return out


@pytest.mark.parametrize("name", DISTORTED)
def test_undistort_restatement_equals_cv2_fixture(name):
    p = random_points(name, 4000, 5)
    assert _undistort(name, p).tobytes() == REF["cv2/" + name].tobytes()


@pytest.mark.parametrize("name", DISTORTED)
def test_undistort_restatement_equals_live_cv2(name):
    cv2 = pytest.importorskip("cv2")
    w, h, cam, dist, bf = CAMERAS[name]
    K = np.array([[cam[0], 0, cam[2]], [0, cam[1], cam[3]], [0, 0, 1]], np.float32)
    p = random_points(name, 20000, 99)
    want = cv2.undistortPoints(p.reshape(-1, 1, 2), K, np.array(dist, np.float32).reshape(-1, 1), None, K).reshape(-1, 2)
    assert _undistort(name, p).tobytes() == want.tobytes()


def _case(case):
    camname, n, seed, with_depth = FINISH_CASES[case]
    x, y = keypoints(camname, n, seed)
    return CAMERAS[camname], x, y, depth_image(camname, seed) if with_depth else None


def _same(got, case):
    for k in ("bounds", "xy_un", "cell_off", "cell_idx", "u_right", "depth"):
        key = "ref/%s/%s" % (case, k)
        if key in REF.files:
            assert got[k].tobytes() == REF[key].tobytes(), (case, k)
        else:
            assert k not in got


@pytest.mark.parametrize("case", sorted(FINISH_CASES))
def test_restatement_reproduces_reference_frame_fixture(case):
    cam, x, y, depth = _case(case)
    _same(oracle_lib.port_frame_finish(oracle(), x, y, cam, depth), case)


@pytest.mark.parametrize("case", sorted(FINISH_CASES))
def test_live_reference_reproduces_fixture(case):
    FR = oracle_lib.frame_ref_lib()
    if FR is None:
        pytest.skip("oracle/_ref/libframeref.so not built (needs /root/reference)")
    cam, x, y, depth = _case(case)
    _same(oracle_lib.ref_frame_finish(FR, x, y, cam, depth), case)


def test_fixture_exercises_the_interesting_paths():
    # keypoints that leave the grid, RGB-D holes, the icdist < 0 escape (bounds fall back to the image)
    assert REF["ref/tum1_rgbd/cell_off"][-1] < 1000
    ur = REF["ref/tum1_rgbd/u_right"]
    assert (ur == -1).sum() > 50 and (ur > 0).sum() > 500
    assert REF["ref/wide_mono/bounds"][:4].tolist() == [0.0, 640.0, 0.0, 480.0]
    assert REF["ref/wide_mono/cell_off"][-1] < 2000


def test_c_abi_fails_loudly_without_a_device():
    import ctypes as C
    from orb_slam2_e_b200 import _lib, frame
    L = _lib.load()
    if L.orb_device_count() > 0:
        pytest.skip("a CUDA device is present")
    cam = frame.camera(500, 500, 320, 240, 0.1)
    with pytest.raises(_lib.OrbFrontError) as e:
        frame.undistort_points([[1.0, 2.0]], cam)
    assert e.value.code == _lib.ORB_E_NO_DEVICE
    with pytest.raises(_lib.OrbFrontError) as e:
        frame.frame_finish(np.zeros(4, _lib.KEYPOINT_DTYPE), cam, frame.frame_bounds(0, 640, 0, 480))
    assert e.value.code == _lib.ORB_E_NO_DEVICE
