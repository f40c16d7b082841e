"""Validates the CPU restatement (oracle/orb_port.hpp) against the reference's own
src/ORBextractor.cc compiled unmodified over oracle/cvshim (oracle/_ref/liborbref.so).
Skipped where oracle/_ref is absent and cannot be built (no /root/reference)."""
import numpy as np
import pytest

from oracle_lib import PortExtractor, RefExtractor, ref_lib
from orb_slam2_e_b200 import synth

pytestmark = pytest.mark.skipif(ref_lib() is None, reason="oracle/_ref not built and /root/reference absent")

CONFIGS = {
    "tum": (1000, 1.2, 8, 20, 7),
    "kitti": (2000, 1.2, 8, 20, 7),
    "hamlyn": (1200, 1.1, 6, 24, 7),
}


def _check(img, cfg):
    ref = RefExtractor(*cfg)
    port = PortExtractor(*cfg)
    rk, rd = ref.extract(img)
    pk, pd = port.extract(img)
    for l in range(cfg[2]):
        assert ref.quota(l) == port.quota(l)
        assert ref.scale(l) == port.scale(l)
        assert np.array_equal(ref.level(l, True), port.level(l, True)), f"pyramid level {l}"
    for v in range(16):
        assert ref.umax(v) == port.umax(v)
    assert len(rk) == len(pk) and len(rk) > 0
    assert rk.tobytes() == pk.tobytes()  # same keypoints, same order, bit-identical floats
    assert np.array_equal(rd, pd)
    return len(rk)


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_textured_tum(seed):
    n = _check(synth.textured_frame(480, 640, seed), CONFIGS["tum"])
    assert n >= 1000


@pytest.mark.parametrize("seed", [11, 12])
def test_endoscopy_hamlyn(seed):
    _check(synth.endoscopy_frame(480, 640, seed), CONFIGS["hamlyn"])


def test_kitti_shape():
    left, right = synth.stereo_pair(376, 1241, 21)
    _check(left, CONFIGS["kitti"])
    _check(right, CONFIGS["kitti"])


def test_small_and_flat():
    _check(synth.textured_frame(120, 160, 5), (300, 1.2, 4, 20, 7))
    ref = RefExtractor(500, 1.2, 3, 20, 7)
    port = PortExtractor(500, 1.2, 3, 20, 7)
    flat = np.full((200, 260), 90, np.uint8)
    rk, rd = ref.extract(flat)
    pk, pd = port.extract(flat)
    assert len(rk) == 0 and len(pk) == 0
