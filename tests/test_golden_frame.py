"""Golden fixture tests/golden/frame_ref.npz (made by tests/golden/make_golden_frame.py from the
reference's own src/Frame.cc, src/MapPoint.cc and src/ORBextractor.cc): the restatement (CPU test) and
the CUDA path (gpu test) must reproduce the reference's mvuRight / mvDepth and distinctive
descriptors.  Neither needs /root/reference at run time."""
import hashlib
import os
import sys

import numpy as np
import pytest

import match_cases as mc
from oracle_lib import PortExtractor, oracle

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
from make_golden_frame_cases import STEREO_CASES, distinctive_case, stereo_pair  # noqa: E402

REF = np.load(os.path.join(HERE, "golden", "frame_ref.npz"))


def _pair(name):
    shape, seed, cfg, max_disp, mb, mbf = STEREO_CASES[name]
    left, right = stereo_pair(shape, seed, max_disp)
    sha = np.frombuffer(hashlib.sha1(left.tobytes() + right.tobytes()).digest(), np.uint8)
    assert np.array_equal(sha, REF[name + "/img_sha1"]), "synthetic stereo generator drifted from the fixture"
    return left, right, cfg, mb, mbf


@pytest.mark.parametrize("name", sorted(STEREO_CASES))
def test_oracle_reproduces_reference_stereo_fixture(name):
    left, right, cfg, mb, mbf = _pair(name)
    pl, pr = PortExtractor(*cfg), PortExtractor(*cfg)
    pl.cap = pr.cap = cfg[0] + 400
    kl, dl = pl.extract(left)
    kr, dr = pr.extract(right)
    assert kl.tobytes() == REF[name + "/kps_left"].tobytes()
    scale = np.array([pl.scale(l) for l in range(cfg[2])], np.float32)
    ur, z = mc.oracle_stereo_matches(oracle(), kl, dl, kr, dr, scale, (np.float32(1.0) / scale).astype(np.float32),
                                     [pl.level(l) for l in range(cfg[2])], [pr.level(l) for l in range(cfg[2])], mb, mbf)
    assert ur.tobytes() == REF[name + "/u_right"].tobytes() and z.tobytes() == REF[name + "/depth"].tobytes()


def test_oracle_reproduces_reference_distinctive_fixture():
    desc, offset = distinctive_case()
    best = np.zeros(len(offset) - 1, np.int32)
    oracle().orc_distinctive_descriptors(desc, offset, len(best), best)
    assert np.array_equal(desc[offset[:-1] + best], REF["distinctive/chosen"])


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(STEREO_CASES))
def test_cuda_reproduces_reference_stereo_fixture(name):
    """Both images extracted on the GPU, ComputeStereoMatches on the resident pyramids: the reference's
    own Frame.cc output, bit for bit (keypoint angles aside, the extractor fields are equal too)."""
    from orb_slam2_e_b200 import extractor as ex
    left, right, cfg, mb, mbf = _pair(name)
    gl, gr = ex.ORBextractor(*cfg), ex.ORBextractor(*cfg)
    kl, dl = gl(left)
    kr, dr = gr(right)
    rk = REF[name + "/kps_left"]
    assert len(kl) == len(rk)
    for f in ("x", "y", "octave", "response"):
        assert np.array_equal(kl[f], rk[f]), f
    ur, z = ex.stereo_matches(gl, gr, kl, dl, kr, dr, mb, mbf)
    assert (REF[name + "/u_right"] >= 0).sum() > 100
    assert ur.tobytes() == REF[name + "/u_right"].tobytes() and z.tobytes() == REF[name + "/depth"].tobytes()


@pytest.mark.gpu
def test_cuda_reproduces_reference_distinctive_fixture():
    from orb_slam2_e_b200 import matcher
    desc, offset = distinctive_case()
    best = matcher.distinctive_descriptors(desc, offset)
    assert np.array_equal(desc[offset[:-1] + best], REF["distinctive/chosen"])
