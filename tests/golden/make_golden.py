"""Generates tests/golden/*.npz IN THE BUILD CONTAINER (needs cv2 4.13 and /root/reference via
oracle/_ref).  The fixtures are outputs of the real things -- cv2's primitives and the reference's
own ORBextractor.cc -- so the GPU box (no cv2 guarantee, no /root/reference) can check both the
oracle restatement and the CUDA path against them.

    python tests/golden/make_golden.py
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import cv2  # noqa: E402
import oracle_lib  # noqa: E402
from orb_slam2_e_b200 import synth  # noqa: E402

from make_golden_cases import EXTRACT_CASES, frame  # noqa: E402


def main():
    assert cv2.__version__.startswith("4.13"), cv2.__version__
    assert oracle_lib.ref_lib() is not None, "oracle/_ref must be built (needs /root/reference)"
    out = {}
    for name, (kind, shape, seed, cfg) in EXTRACT_CASES.items():
        img = frame(kind, shape, seed)
        ref = oracle_lib.RefExtractor(*cfg)
        k, d = ref.extract(img)
        out[name + "/kps"] = k
        out[name + "/desc"] = d
        out[name + "/img_sha1"] = np.frombuffer(hashlib.sha1(img.tobytes()).digest(), np.uint8)
        # last pyramid level with its border: a compact checksum-like pin of the whole resize chain
        out[name + "/last_level"] = ref.level(cfg[2] - 1, True)
        print(name, len(k), "keypoints")
    np.savez_compressed(os.path.join(HERE, "extract_ref.npz"), **out)

    # cv2 primitives on a small frame
    img = synth.textured_frame(96, 128, 3)
    prim = {"img": img,
            "resize": cv2.resize(img, (107, 80), interpolation=cv2.INTER_LINEAR),
            "border": cv2.copyMakeBorder(img, 19, 19, 19, 19, cv2.BORDER_REFLECT_101),
            "blur": cv2.GaussianBlur(img, (7, 7), 2, sigmaY=2, borderType=cv2.BORDER_REFLECT_101)}
    for th in (20, 7):
        kp = cv2.FastFeatureDetector_create(th, True, cv2.FAST_FEATURE_DETECTOR_TYPE_9_16).detect(img)
        prim[f"fast{th}"] = np.array([(int(p.pt[0]), int(p.pt[1]), int(p.response)) for p in kp], np.int32).reshape(-1, 3)
    rng = np.random.default_rng(5)
    y = rng.integers(-200000, 200000, 4000).astype(np.float32)
    x = rng.integers(-200000, 200000, 4000).astype(np.float32)
    prim["atan_y"], prim["atan_x"] = y, x
    prim["atan"] = np.array([cv2.fastAtan2(float(a), float(b)) for a, b in zip(y, x)], np.float32)
    np.savez_compressed(os.path.join(HERE, "cv2_prims.npz"), **prim)
    for f in ("extract_ref.npz", "cv2_prims.npz"):
        print(f, os.path.getsize(os.path.join(HERE, f)), "bytes")


if __name__ == "__main__":
    main()
