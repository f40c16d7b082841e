"""Golden fixtures (tests/golden/*.npz, made by tests/golden/make_golden.py from cv2 4.13 and from
the reference's own ORBextractor.cc): the oracle restatement must reproduce them on any box (CPU
tests), and so must the CUDA path (gpu tests) -- neither needs cv2 or /root/reference at run time."""
import hashlib
import os
import sys

import numpy as np
import pytest

from oracle_lib import PortExtractor, oracle
from orb_slam2_e_b200 import synth

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
from make_golden_cases import EXTRACT_CASES, frame  # noqa: E402

REF = np.load(os.path.join(HERE, "golden", "extract_ref.npz"))
PRIM = np.load(os.path.join(HERE, "golden", "cv2_prims.npz"))


def _golden_frame(name):
    kind, shape, seed, cfg = EXTRACT_CASES[name]
    img = frame(kind, shape, seed)
    sha = np.frombuffer(hashlib.sha1(img.tobytes()).digest(), np.uint8)
    assert np.array_equal(sha, REF[name + "/img_sha1"]), "synthetic frame generator drifted from the fixture"
    return img, cfg


def _check_against_reference(name, kps, desc, last_level):
    rk, rd = REF[name + "/kps"], REF[name + "/desc"]
    assert len(kps) == len(rk)
    for f in ("x", "y", "size", "response", "octave", "class_id"):
        assert np.array_equal(kps[f], rk[f]), f
    dang = np.abs(np.deg2rad(kps["angle"].astype(np.float64) - rk["angle"]))
    assert np.minimum(dang, 2 * np.pi - dang).max(initial=0) <= 1e-3
    assert np.unpackbits(desc ^ rd).mean() <= 1e-4
    assert np.array_equal(last_level, REF[name + "/last_level"])


@pytest.mark.parametrize("name", sorted(EXTRACT_CASES))
def test_oracle_reproduces_reference_fixture(name):
    img, cfg = _golden_frame(name)
    p = PortExtractor(*cfg)
    k, d = p.extract(img)
    assert k.tobytes() == REF[name + "/kps"].tobytes() and np.array_equal(d, REF[name + "/desc"])
    assert np.array_equal(p.level(cfg[2] - 1, True), REF[name + "/last_level"])


def test_oracle_reproduces_cv2_fixture():
    L = oracle()
    img = PRIM["img"]
    h, w = img.shape
    dst = np.zeros_like(PRIM["resize"])
    L.orc_resize_linear(img, w, h, img.strides[0], dst, dst.shape[1], dst.shape[0], dst.strides[0])
    assert np.array_equal(dst, PRIM["resize"])
    dst = np.zeros_like(PRIM["border"])
    L.orc_border101(img, w, h, img.strides[0], dst, dst.strides[0], 19)
    assert np.array_equal(dst, PRIM["border"])
    dst = np.zeros_like(img)
    L.orc_blur7(img, w, h, img.strides[0], dst, dst.strides[0])
    assert np.array_equal(dst, PRIM["blur"])
    for th in (20, 7):
        buf = np.zeros((h * w, 3), np.int32)
        n = L.orc_fast(img.ctypes.data, w, h, img.strides[0], th, buf.reshape(-1), len(buf))
        assert np.array_equal(buf[:n], PRIM[f"fast{th}"]) and n > 20
    out = np.zeros_like(PRIM["atan"])
    L.orc_fast_atan2_v(PRIM["atan_y"], PRIM["atan_x"], out, len(out))
    assert np.array_equal(out, PRIM["atan"])


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(EXTRACT_CASES))
def test_cuda_reproduces_reference_fixture(name):
    from orb_slam2_e_b200.extractor import ORBextractor
    img, cfg = _golden_frame(name)
    g = ORBextractor(*cfg)
    k, d = g(img)
    _check_against_reference(name, k, d, g.level(cfg[2] - 1, 0, True))
