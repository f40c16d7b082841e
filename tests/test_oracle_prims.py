"""Pins the oracle's OpenCV-primitive restatements (oracle/cvprims.hpp) against cv2 4.13.

OpenCV is not vendored in the reference (CMakeLists.txt:20), so these known-answer checks are
what "bit-exact with the reference" rests on for resize / copyMakeBorder / FAST / GaussianBlur /
fastAtan2 (reference call sites: src/ORBextractor.cc:103, :810, :822, :1094, :1128-1136)."""
import ctypes as C
import numpy as np
import pytest

from oracle_lib import oracle
from orb_slam2_e_b200 import synth

cv2 = pytest.importorskip("cv2")


def _images():
    rng = np.random.default_rng(7)
    yield "random", rng.integers(0, 256, (97, 131), dtype=np.uint8)
    yield "textured", synth.textured_frame(120, 160, 3)
    yield "endo", synth.endoscopy_frame(120, 160, 4)
    yield "const255", np.full((40, 50), 255, np.uint8)
    q = (rng.integers(0, 4, (80, 90)) * 60).astype(np.uint8)  # tie-heavy
    yield "quantised", q


@pytest.mark.parametrize("shape,dshape", [((480, 640), (400, 533)), ((376, 1241), (313, 1034)),
                                          ((400, 533), (333, 444)), ((97, 131), (81, 109)), ((50, 60), (50, 60))])
def test_resize_matches_cv2(shape, dshape):
    rng = np.random.default_rng(shape[0])
    src = rng.integers(0, 256, shape, dtype=np.uint8)
    dst = np.zeros(dshape, np.uint8)
    oracle().orc_resize_linear(src, shape[1], shape[0], src.strides[0], dst, dshape[1], dshape[0], dst.strides[0])
    want = cv2.resize(src, (dshape[1], dshape[0]), interpolation=cv2.INTER_LINEAR)
    assert np.array_equal(dst, want)


def test_border_matches_cv2():
    for name, img in _images():
        h, w = img.shape
        dst = np.zeros((h + 38, w + 38), np.uint8)
        oracle().orc_border101(img, w, h, img.strides[0], dst, dst.strides[0], 19)
        want = cv2.copyMakeBorder(img, 19, 19, 19, 19, cv2.BORDER_REFLECT_101)
        assert np.array_equal(dst, want), name


def test_blur_matches_cv2():
    for name, img in _images():
        h, w = img.shape
        dst = np.zeros_like(img)
        oracle().orc_blur7(img, w, h, img.strides[0], dst, dst.strides[0])
        want = cv2.GaussianBlur(img, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)
        assert np.array_equal(dst, want), name


def test_fast_atan2_matches_cv2():
    rng = np.random.default_rng(5)
    y = rng.integers(-200000, 200000, 20000).astype(np.float32)
    x = rng.integers(-200000, 200000, 20000).astype(np.float32)
    y[:50] = 0
    x[25:75] = 0
    out = np.zeros_like(x)
    oracle().orc_fast_atan2_v(y, x, out, len(x))
    want = np.array([cv2.fastAtan2(float(a), float(b)) for a, b in zip(y, x)], np.float32)
    assert np.array_equal(out, want)
    assert np.max(np.abs(np.deg2rad(out[75:]) - np.mod(np.arctan2(y[75:], x[75:]), 2 * np.pi))) < 1e-3


@pytest.mark.parametrize("th", [20, 7, 24])
def test_fast_matches_cv2(th):
    det = cv2.FastFeatureDetector_create(threshold=th, nonmaxSuppression=True,
                                         type=cv2.FastFeatureDetector_TYPE_9_16)
    for name, img in _images():
        # also a non-contiguous view, as the reference passes cell ROIs
        for view in (img, img[5:45, 7:50]):
            kps = det.detect(np.ascontiguousarray(view) if False else view)
            want = [(int(k.pt[0]), int(k.pt[1]), int(k.response)) for k in kps]
            buf = np.zeros((view.shape[0] * view.shape[1], 3), np.int32)
            n = oracle().orc_fast(view.ctypes.data, view.shape[1], view.shape[0], view.strides[0], th,
                                  buf.reshape(-1), len(buf))
            got = [tuple(r) for r in buf[:n].tolist()]
            assert got == want, (name, th, len(got), len(want))
