"""Generates tests/golden/frame_ref.npz IN THE BUILD CONTAINER from the reference's own src/Frame.cc
and src/MapPoint.cc (oracle/_ref/libframeref.so, needs /root/reference): mvuRight / mvDepth of
Frame::ComputeStereoMatches on seeded stereo pairs and the descriptors chosen by
MapPoint::ComputeDistinctiveDescriptors.  The inputs are regenerated from seeds by the tests, so the
GPU box needs neither the reference nor these libraries.

    python tests/golden/make_golden_frame.py
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import oracle_lib  # noqa: E402
from make_golden_frame_cases import STEREO_CASES, distinctive_case, stereo_pair  # noqa: E402


def main():
    FR = oracle_lib.frame_ref_lib()
    assert FR is not None and oracle_lib.ref_lib() is not None, "oracle/_ref must be built (needs /root/reference)"
    out = {}
    for name, (shape, seed, cfg, max_disp, mb, mbf) in STEREO_CASES.items():
        left, right = stereo_pair(shape, seed, max_disp)
        rl, rr = oracle_lib.RefExtractor(*cfg), oracle_lib.RefExtractor(*cfg)  # the reference's own extractor
        rl.cap = rr.cap = cfg[0] + 400
        kl, dl = rl.extract(left)
        kr, dr = rr.extract(right)
        ur, z = oracle_lib.ref_stereo_matches(FR, cfg, kl, dl, kr, dr, [rl.level(l) for l in range(cfg[2])],
                                              [rr.level(l) for l in range(cfg[2])], mb, mbf)
        out[name + "/img_sha1"] = np.frombuffer(hashlib.sha1(left.tobytes() + right.tobytes()).digest(), np.uint8)
        out[name + "/kps_left"] = kl
        out[name + "/u_right"] = ur
        out[name + "/depth"] = z
        print(name, len(kl), "left keypoints,", int((ur >= 0).sum()), "stereo matches")
    desc, offset = distinctive_case()
    chosen = np.zeros((len(offset) - 1, 32), np.uint8)
    FR.frameref_distinctive(desc, offset, None, len(offset) - 1, chosen)
    out["distinctive/chosen"] = chosen
    np.savez_compressed(os.path.join(HERE, "frame_ref.npz"), **out)
    print("frame_ref.npz", os.path.getsize(os.path.join(HERE, "frame_ref.npz")), "bytes")


if __name__ == "__main__":
    main()
